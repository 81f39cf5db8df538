from .caffe_loader import CaffeConversionException, CaffeGraph, CaffeLoader, load_caffe, save_caffemodel  # noqa: F401
