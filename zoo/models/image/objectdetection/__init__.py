from .object_detector import (DecodeOutput, ImInfo, ObjectDetector, ScaleDetection, Visualizer, read_coco_label_map,  # noqa: F401
                              read_pascal_label_map)
