"""TFPark text estimators (reference: pyzoo/zoo/tfpark/text/**: BERT classifier / NER / SQuAD estimators and the
NLP-architect keras models NER, SequenceTagger, IntentEntity).  They are TensorFlow model functions; the torch
counterparts are `zoo.pipeline.api.keras.layers.BERT` / `TransformerLayer` with a task head trained by `Estimator`."""
from ..unsupported import _stub

BERTClassifier = _stub("BERTClassifier")
BERTNER = _stub("BERTNER")
BERTSQuAD = _stub("BERTSQuAD")
NER = _stub("NER")
SequenceTagger = _stub("SequenceTagger")
IntentEntity = _stub("IntentEntity")
