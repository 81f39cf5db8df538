"""TFPark text models (reference: pyzoo/zoo/tfpark/text/**): the BERT estimators and the Keras-style NER / SequenceTagger /
IntentEntity models, as torch networks with the reference's constructor arguments and train / evaluate / predict surface."""
from .estimator import BERTNER, BERTBaseEstimator, BERTClassifier, BERTSQuAD, BertEncoder, bert_input_fn  # noqa: F401
from .keras import CRF, NER, IntentEntity, SequenceTagger, TextKerasModel  # noqa: F401
