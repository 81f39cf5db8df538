"""Keras-2 named layers (reference: pyzoo/zoo/pipeline/api/keras2/layers/*, Scala keras2/layers/*.scala): thin
re-namings of the keras-1 layers with Keras 2 argument names (`units`, `filters`, `kernel_size`, `strides`,
`padding`, `kernel_initializer`, `use_bias`)."""
from __future__ import annotations

from ...keras import layers as k1
from ...keras.layers.core import Merge


def _init(kernel_initializer):
    return kernel_initializer or "glorot_uniform"


class Dense(k1.Dense):
    def __init__(self, units, kernel_initializer="glorot_uniform", bias_initializer="zero", activation=None, kernel_regularizer=None,
                 bias_regularizer=None, use_bias=True, input_dim=None, input_shape=None, **kw):
        super().__init__(units, _init(kernel_initializer), activation, kernel_regularizer, bias_regularizer, use_bias, input_dim, input_shape, **kw)


class Activation(k1.Activation):
    pass


class Dropout(k1.Dropout):
    def __init__(self, rate, input_shape=None, **kw):
        super().__init__(rate, input_shape, **kw)


class Flatten(k1.Flatten):
    pass


class Conv1D(k1.Convolution1D):
    def __init__(self, filters, kernel_size, strides=1, padding="valid", activation=None, use_bias=True,
                 kernel_initializer="glorot_uniform", bias_initializer="zero", kernel_regularizer=None, bias_regularizer=None,
                 input_shape=None, **kw):
        super().__init__(filters, kernel_size, _init(kernel_initializer), activation, padding, strides, kernel_regularizer,
                         bias_regularizer, use_bias, input_shape, **kw)


class Conv2D(k1.Convolution2D):
    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", data_format="channels_first", activation=None,
                 use_bias=True, kernel_initializer="glorot_uniform", bias_initializer="zero", kernel_regularizer=None,
                 bias_regularizer=None, input_shape=None, **kw):
        ks = (kernel_size, kernel_size) if isinstance(kernel_size, int) else tuple(kernel_size)
        st = (strides, strides) if isinstance(strides, int) else tuple(strides)
        super().__init__(filters, ks[0], ks[1], _init(kernel_initializer), activation, padding, st,
                         "th" if data_format == "channels_first" else "tf", kernel_regularizer, bias_regularizer, use_bias, input_shape, **kw)


class Cropping1D(k1.Cropping1D):
    pass


class _Pool2Mixin:
    @staticmethod
    def _fmt(data_format):
        return "th" if data_format == "channels_first" else "tf"


class MaxPooling1D(k1.MaxPooling1D):
    def __init__(self, pool_size=2, strides=None, padding="valid", input_shape=None, **kw):
        super().__init__(pool_size, strides, padding, input_shape, **kw)


class AveragePooling1D(k1.AveragePooling1D):
    def __init__(self, pool_size=2, strides=None, padding="valid", input_shape=None, **kw):
        super().__init__(pool_size, strides, padding, input_shape, **kw)


class MaxPooling2D(k1.MaxPooling2D, _Pool2Mixin):
    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", data_format="channels_first", input_shape=None, **kw):
        super().__init__(pool_size, strides, padding, self._fmt(data_format), input_shape, **kw)


class AveragePooling2D(k1.AveragePooling2D, _Pool2Mixin):
    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", data_format="channels_first", input_shape=None, **kw):
        super().__init__(pool_size, strides, padding, self._fmt(data_format), input_shape, **kw)


class GlobalMaxPooling1D(k1.GlobalMaxPooling1D):
    pass


class GlobalAveragePooling1D(k1.GlobalAveragePooling1D):
    pass


class GlobalMaxPooling2D(k1.GlobalMaxPooling2D, _Pool2Mixin):
    def __init__(self, data_format="channels_first", input_shape=None, **kw):
        super().__init__(self._fmt(data_format), input_shape, **kw)


class GlobalAveragePooling2D(k1.GlobalAveragePooling2D, _Pool2Mixin):
    def __init__(self, data_format="channels_first", input_shape=None, **kw):
        super().__init__(self._fmt(data_format), input_shape, **kw)


class GlobalMaxPooling3D(k1.GlobalMaxPooling3D, _Pool2Mixin):
    def __init__(self, data_format="channels_first", input_shape=None, **kw):
        super().__init__(self._fmt(data_format), input_shape, **kw)


class GlobalAveragePooling3D(k1.GlobalAveragePooling3D, _Pool2Mixin):
    def __init__(self, data_format="channels_first", input_shape=None, **kw):
        super().__init__(self._fmt(data_format), input_shape, **kw)


class LocallyConnected1D(k1.LocallyConnected1D):
    def __init__(self, filters, kernel_size, strides=1, padding="valid", activation=None, kernel_regularizer=None,
                 bias_regularizer=None, use_bias=True, input_shape=None, **kw):
        super().__init__(filters, kernel_size, activation, padding, strides, kernel_regularizer, bias_regularizer, use_bias, input_shape, **kw)


class _MergeMode(Merge):
    mode_name = "sum"

    def __init__(self, input_shape=None, **kw):
        super().__init__(None, self.mode_name, -1, input_shape, **kw)


class Maximum(_MergeMode):
    mode_name = "max"


class Minimum(_MergeMode):
    mode_name = "min"


class Average(_MergeMode):
    mode_name = "ave"


class Add(_MergeMode):
    mode_name = "sum"


class Multiply(_MergeMode):
    mode_name = "mul"


class Concatenate(Merge):
    def __init__(self, axis=-1, input_shape=None, **kw):
        super().__init__(None, "concat", axis, input_shape, **kw)


def maximum(inputs, **kw):
    return Maximum(**kw)(list(inputs))


def minimum(inputs, **kw):
    return Minimum(**kw)(list(inputs))


def average(inputs, **kw):
    return Average(**kw)(list(inputs))
