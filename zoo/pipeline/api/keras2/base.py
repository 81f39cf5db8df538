"""Keras-2 layer base names (reference: pyzoo/zoo/pipeline/api/keras2/base.py)."""
from ..keras.layers.base import ZooKerasLayer

ZooKeras2Layer = ZooKerasLayer
ZooKeras2Creator = ZooKerasLayer
