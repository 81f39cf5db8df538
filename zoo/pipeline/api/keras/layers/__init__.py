"""zoo.pipeline.api.keras.layers: the keras-style layer library (reference: pyzoo/zoo/pipeline/api/keras/layers/*)."""
from .advanced_activations import *  # noqa: F401,F403
from .advanced_activations import ELU, LeakyReLU, SReLU, ThresholdedReLU  # noqa: F401
from .base import Input, InputLayer, ZooKerasLayer  # noqa: F401
from .convolutional import (AtrousConvolution1D, AtrousConvolution2D, Convolution1D, Convolution2D, Convolution3D,  # noqa: F401
                            Cropping1D, Cropping2D, Cropping3D, Deconvolution2D, SeparableConvolution2D, UpSampling1D,
                            UpSampling2D, UpSampling3D, ZeroPadding1D, ZeroPadding2D, ZeroPadding3D)
from .convolutional_recurrent import ConvLSTM2D, ConvLSTM3D  # noqa: F401
from .core import (Activation, Dense, Dropout, ExpandDim, Flatten, GetShape, Highway, Masking, Max, MaxoutDense, Merge,  # noqa: F401
                   Permute, RepeatVector, Reshape, SparseDense, SpatialDropout1D, SpatialDropout2D, SpatialDropout3D, merge)
from .embeddings import Embedding, SparseEmbedding, WordEmbedding  # noqa: F401
from .local import LocallyConnected1D, LocallyConnected2D  # noqa: F401
from .noise import GaussianDropout, GaussianNoise  # noqa: F401
from .normalization import BatchNormalization, LayerNorm  # noqa: F401
from .pooling import (AveragePooling1D, AveragePooling2D, AveragePooling3D, GlobalAveragePooling1D, GlobalAveragePooling2D,  # noqa: F401
                      GlobalAveragePooling3D, GlobalMaxPooling1D, GlobalMaxPooling2D, GlobalMaxPooling3D, MaxPooling1D,
                      MaxPooling2D, MaxPooling3D)
from .recurrent import GRU, LSTM, SimpleRNN  # noqa: F401
from .self_attention import BERT, TransformerLayer  # noqa: F401
from .torch import (AddConstant, BinaryThreshold, CAdd, CMul, Exp, Expand, GaussianSampler, HardShrink, HardTanh, Identity,  # noqa: F401
                    Log, LRN2D, Mul, MulConstant, Narrow, Negative, Power, PReLU, ResizeBilinear, RReLU, Scale, Select,
                    SelectTable, ShareConvolution2D, SoftShrink, Softmax, SplitTensor, Sqrt, Square, Squeeze, Threshold,
                    WithinChannelLRN2D)
from .wrappers import Bidirectional, KerasLayerWrapper, TimeDistributed  # noqa: F401

Conv1D, Conv2D, Conv3D = Convolution1D, Convolution2D, Convolution3D
