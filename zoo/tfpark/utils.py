"""`zoo.tfpark.utils` (reference: pyzoo/zoo/tfpark/utils.py: evaluate_string_metrics / evaluate_metrics over a TFNet)."""
import numpy as np


def evaluate_string_metrics(*, sess=None, string_metrics, dataset, inputs=None, targets=None, outputs=None, loss=None, model=None):
    """Evaluate the named metrics of a zoo model over a TFDataset (the TF-session arguments of the reference are ignored)."""
    if model is None:
        raise NotImplementedError("evaluate_string_metrics over a TensorFlow session needs the TensorFlow runtime; pass model=<zoo KerasNet>")
    res = model.evaluate(dataset.get_evaluation_data(), batch_size=max(dataset.batch_size, dataset.batch_per_thread, 1))
    return {m: float(np.asarray(v)) for m, v in zip(string_metrics, res)}


def evaluate_metrics(inputs, sess, dataset, metrics):
    raise NotImplementedError("evaluate_metrics drives a TensorFlow session and needs the TensorFlow runtime, which is not part of zoo-b200")
