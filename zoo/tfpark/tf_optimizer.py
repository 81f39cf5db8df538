"""`zoo.tfpark.tf_optimizer` names (reference: pyzoo/zoo/tfpark/tf_optimizer.py:40-709).  TFOptimizer / TFTrainingHelper / TFModel
run a TensorFlow graph inside the BigDL optimizer and need the TF runtime.  The metric adapters keep their meaning:
`StatelessMetric(name, idx)` averages the idx-th extra output of the model, `BigDLMetric` wraps a zoo validation method."""
from ..pipeline.api.keras import metrics as _metrics
from .unsupported import TFOptimizer, _stub  # noqa: F401

TFTrainingHelper = _stub("TFTrainingHelper")
TFModel = _stub("TFModel", ("create", "create_for_unfreeze"))


class IdentityCriterion:
    """The loss is computed inside the model graph; the criterion returns its first output unchanged."""

    def __call__(self, output, target=None):
        out = output[0] if isinstance(output, (list, tuple)) else output
        return out.mean() if hasattr(out, "mean") else out


class TFValidationMethod:
    """(val_method, name, output_indices, label_indices): picks outputs / labels by index and applies `val_method`."""

    def __init__(self, val_method, name, output_indices, label_indices):
        self.val_method, self.name, self.output_indices, self.label_indices = val_method, name, list(output_indices), list(label_indices)

    def __call__(self, outputs, labels):
        pick = lambda xs, idx: [xs[i] for i in idx] if isinstance(xs, (list, tuple)) else xs
        o, l = pick(outputs, self.output_indices), pick(labels, self.label_indices)
        return self.val_method(o[0] if isinstance(o, list) and len(o) == 1 else o, l[0] if isinstance(l, list) and len(l) == 1 else l)


class StatelessMetric:
    """Mean over batches of the `idx`-th metric tensor the model produces."""

    def __init__(self, metric_name, idx):
        self.name, self.idx, self.total, self.count = metric_name, idx, 0.0, 0

    def update(self, outputs, batch_size=1):
        v = outputs[self.idx] if isinstance(outputs, (list, tuple)) else outputs
        self.total += float(v.mean() if hasattr(v, "mean") else v) * batch_size
        self.count += batch_size
        return self

    def result(self):
        return self.total / max(self.count, 1), self.count


class BigDLMetric:
    def __init__(self, val_method, outputs=None, labels=None):
        self.val_method = _metrics.to_metric(val_method) if hasattr(_metrics, "to_metric") else val_method
        self.outputs, self.labels = outputs, labels
