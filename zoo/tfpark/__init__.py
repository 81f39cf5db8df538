"""TFPark names (reference: pyzoo/zoo/tfpark/*).  TensorFlow is not part of zoo-b200 (dropped as an
engine), so what is kept is what does not need a TF runtime: `TFDataset` as a batched data container that feeds the
training engine, `KerasModel` as the fit / evaluate / predict / save_weights face over a zoo (torch) model, and
`ZooOptimizer`; `GANEstimator` and the text models (`zoo.tfpark.text`: BERT classifier / NER / SQuAD estimators, NER,
SequenceTagger, IntentEntity) are torch implementations with the reference's surface.  `TFOptimizer`, `TFEstimator`, `TFNet`
and `TFPredictor` drive TF graphs and raise `NotImplementedError` naming the torch route."""
from .model import KerasModel  # noqa: F401
from .tf_dataset import TFDataset  # noqa: F401
from .gan import GANEstimator  # noqa: F401
from .unsupported import TFEstimator, TFNet, TFOptimizer, TFPredictor, ZooOptimizer  # noqa: F401
