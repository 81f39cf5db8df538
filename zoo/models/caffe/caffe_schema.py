"""Field numbers of the parts of Caffe's `caffe.proto` the importer reads (NetParameter, LayerParameter, V1LayerParameter, BlobProto
and the per-layer parameter messages), as schemas for :mod:`zoo.util.protowire`.  Names equal the prototxt field names, so a
binary `.caffemodel` and a text `.prototxt` decode into the same dict shape.
Reference: the reference reads the same messages through generated `caffe.Caffe` classes (models/caffe/CaffeLoader.scala:91-150)."""

BLOB_SHAPE = {1: ("dim", "packed_varint")}
BLOB = {1: ("num", "varint"), 2: ("channels", "varint"), 3: ("height", "varint"), 4: ("width", "varint"),
        5: ("data", "packed_float"), 7: ("shape", "msg", BLOB_SHAPE), 8: ("double_data", "packed_double")}
FILLER = {1: ("type", "string"), 2: ("value", "float")}

CONVOLUTION = {1: ("num_output", "varint"), 2: ("bias_term", "bool"), 3: ("pad", "packed_varint"), 4: ("kernel_size", "packed_varint"),
               5: ("group", "varint"), 6: ("stride", "packed_varint"), 9: ("pad_h", "varint"), 10: ("pad_w", "varint"),
               11: ("kernel_h", "varint"), 12: ("kernel_w", "varint"), 13: ("stride_h", "varint"), 14: ("stride_w", "varint"),
               16: ("axis", "varint"), 18: ("dilation", "packed_varint")}
INNER_PRODUCT = {1: ("num_output", "varint"), 2: ("bias_term", "bool"), 5: ("axis", "varint"), 6: ("transpose", "bool")}
POOLING = {1: ("pool", "enum"), 2: ("kernel_size", "varint"), 3: ("stride", "varint"), 4: ("pad", "varint"), 5: ("kernel_h", "varint"),
           6: ("kernel_w", "varint"), 7: ("stride_h", "varint"), 8: ("stride_w", "varint"), 9: ("pad_h", "varint"), 10: ("pad_w", "varint"),
           12: ("global_pooling", "bool")}
LRN = {1: ("local_size", "varint"), 2: ("alpha", "float"), 3: ("beta", "float"), 4: ("norm_region", "enum"), 5: ("k", "float")}
DROPOUT = {1: ("dropout_ratio", "float")}
CONCAT = {1: ("concat_dim", "varint"), 2: ("axis", "varint")}
ELTWISE = {1: ("operation", "enum"), 2: ("coeff", "packed_float")}
BATCH_NORM = {1: ("use_global_stats", "bool"), 2: ("moving_average_fraction", "float"), 3: ("eps", "float")}
SCALE = {1: ("axis", "varint"), 2: ("num_axes", "varint"), 4: ("bias_term", "bool")}
BIAS = {1: ("axis", "varint"), 2: ("num_axes", "varint")}
RELU = {1: ("negative_slope", "float")}
ELU = {1: ("alpha", "float")}
POWER = {1: ("power", "float"), 2: ("scale", "float"), 3: ("shift", "float")}
EXPLOG = {1: ("base", "float"), 2: ("scale", "float"), 3: ("shift", "float")}
SOFTMAX = {2: ("axis", "varint")}
FLATTEN = {1: ("axis", "varint"), 2: ("end_axis", "varint")}
RESHAPE = {1: ("shape", "msg", BLOB_SHAPE), 2: ("axis", "varint"), 3: ("num_axes", "varint")}
SLICE = {1: ("slice_dim", "varint"), 2: ("slice_point", "packed_varint"), 3: ("axis", "varint")}
TILE = {1: ("axis", "varint"), 2: ("tiles", "varint")}
THRESHOLD = {1: ("threshold", "float")}
PRELU = {2: ("channel_shared", "bool")}
INPUT = {1: ("shape", "msg", BLOB_SHAPE)}
NET_STATE_RULE = {1: ("phase", "enum")}

LAYER = {1: ("name", "string"), 2: ("type", "string"), 3: ("bottom", "string"), 4: ("top", "string"), 7: ("blobs", "msg", BLOB),
         8: ("include", "msg", NET_STATE_RULE), 9: ("exclude", "msg", NET_STATE_RULE), 10: ("phase", "enum"),
         104: ("concat_param", "msg", CONCAT), 106: ("convolution_param", "msg", CONVOLUTION), 108: ("dropout_param", "msg", DROPOUT),
         110: ("eltwise_param", "msg", ELTWISE), 111: ("exp_param", "msg", EXPLOG), 117: ("inner_product_param", "msg", INNER_PRODUCT),
         118: ("lrn_param", "msg", LRN), 121: ("pooling_param", "msg", POOLING), 122: ("power_param", "msg", POWER),
         123: ("relu_param", "msg", RELU), 125: ("softmax_param", "msg", SOFTMAX), 126: ("slice_param", "msg", SLICE),
         128: ("threshold_param", "msg", THRESHOLD), 131: ("prelu_param", "msg", PRELU), 133: ("reshape_param", "msg", RESHAPE),
         134: ("log_param", "msg", EXPLOG), 135: ("flatten_param", "msg", FLATTEN), 138: ("tile_param", "msg", TILE),
         139: ("batch_norm_param", "msg", BATCH_NORM), 140: ("elu_param", "msg", ELU), 141: ("bias_param", "msg", BIAS),
         142: ("scale_param", "msg", SCALE), 143: ("input_param", "msg", INPUT)}

V1_TYPES = {0: "NONE", 35: "ABSVAL", 1: "ACCURACY", 30: "ARGMAX", 2: "BNLL", 3: "CONCAT", 37: "CONTRASTIVE_LOSS", 4: "CONVOLUTION",
            5: "DATA", 39: "DECONVOLUTION", 6: "DROPOUT", 32: "DUMMY_DATA", 7: "EUCLIDEAN_LOSS", 25: "ELTWISE", 38: "EXP", 8: "FLATTEN",
            9: "HDF5_DATA", 10: "HDF5_OUTPUT", 28: "HINGE_LOSS", 11: "IM2COL", 12: "IMAGE_DATA", 13: "INFOGAIN_LOSS", 14: "INNER_PRODUCT",
            15: "LRN", 29: "MEMORY_DATA", 16: "MULTINOMIAL_LOGISTIC_LOSS", 34: "MVN", 17: "POOLING", 26: "POWER", 18: "RELU",
            19: "SIGMOID", 27: "SIGMOID_CROSS_ENTROPY_LOSS", 36: "SILENCE", 20: "SOFTMAX", 21: "SOFTMAX_LOSS", 22: "SPLIT", 33: "SLICE",
            23: "TANH", 24: "WINDOW_DATA", 31: "THRESHOLD"}
V1_LAYER = {2: ("bottom", "string"), 3: ("top", "string"), 4: ("name", "string"), 5: ("type", "enum"), 6: ("blobs", "msg", BLOB),
            32: ("include", "msg", NET_STATE_RULE), 33: ("exclude", "msg", NET_STATE_RULE),
            9: ("concat_param", "msg", CONCAT), 10: ("convolution_param", "msg", CONVOLUTION), 12: ("dropout_param", "msg", DROPOUT),
            24: ("eltwise_param", "msg", ELTWISE), 41: ("exp_param", "msg", EXPLOG), 17: ("inner_product_param", "msg", INNER_PRODUCT),
            18: ("lrn_param", "msg", LRN), 19: ("pooling_param", "msg", POOLING), 21: ("power_param", "msg", POWER),
            30: ("relu_param", "msg", RELU), 39: ("softmax_param", "msg", SOFTMAX), 31: ("slice_param", "msg", SLICE),
            25: ("threshold_param", "msg", THRESHOLD)}

NET = {1: ("name", "string"), 2: ("layers", "msg", V1_LAYER), 3: ("input", "string"), 4: ("input_dim", "packed_varint"),
       8: ("input_shape", "msg", BLOB_SHAPE), 100: ("layer", "msg", LAYER)}
