from .image_classification import ImageClassifier, LabelOutput, read_imagenet_label_map  # noqa: F401
