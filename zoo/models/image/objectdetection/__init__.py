from .object_detector import (DecodeOutput, ImInfo, ObjectDetector, ScaleDetection, Visualizer, read_coco_label_map,  # noqa: F401
                              read_pascal_label_map)
from .dataset import (Coco, CustomizedDataSet, Imdb, PascalVoc, RoiByteImageToSeq, RoiImageSeqGenerator, SSDDataSet,  # noqa: F401,E402
                      load_roi_seq_files, roi_batch)
from .evaluation import MeanAveragePrecision, PascalVocEvaluator, voc_ap  # noqa: F401,E402
from .ssd import (SSD, SSDVGG, DetectionOutput, MultiBoxLoss, MultiBoxLossParam, PriorBoxSpec, decode_boxes, encode_boxes,  # noqa: F401,E402
                  jaccard, match_priors, nms, prior_boxes, ssd_specs)
