"""`from zoo.pipeline.api.keras.models import Sequential, Model` (reference: pyzoo/zoo/pipeline/api/keras/models.py)."""
from .engine.topology import KerasNet, Model, Sequential  # noqa: F401
