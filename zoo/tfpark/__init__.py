"""TFPark (reference: pyzoo/zoo/tfpark/*).  TensorFlow is not part of zoo-b200 (dropped as an engine);
every TFPark name exists with torch content and the reference's surface: `TFDataset` (batched data container feeding the
training engine; TFRecord files read natively), `KerasModel` (fit / evaluate / predict / save_weights over a zoo model),
`TFOptimizer` (from_loss / from_keras / optimize), `TFEstimator` (from_model_fn with a torch `TFEstimatorSpec`), `TFNet` (frozen
GraphDef imported into torch), `TFPredictor`, `ZooOptimizer`, `GANEstimator` and the text models (`zoo.tfpark.text`: BERT
classifier / NER / SQuAD estimators, NER, SequenceTagger, IntentEntity).  What stays TF-only is what takes a live TensorFlow
object: `TFOptimizer.from_train_op`, `TFDataset.from_tf_data_dataset`, `TFNet.from_session`."""
from .model import KerasModel  # noqa: F401
from .tf_dataset import TFDataset  # noqa: F401
from .gan import GANEstimator  # noqa: F401
from .unsupported import TFNet, ZooOptimizer  # noqa: F401
from .tf_optimizer import TFOptimizer  # noqa: F401
from .estimator import TFEstimator, TFEstimatorSpec  # noqa: F401
from .tf_predictor import TFPredictor  # noqa: F401
