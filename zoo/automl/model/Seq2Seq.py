"""`LSTMSeq2Seq` of the reference is a tf.keras encoder/decoder (pyzoo/zoo/automl/model/Seq2Seq.py:30-330: config keys
latent_dim / dropout / lr / batch_size / epochs, teacher forcing in fit, step-wise decoding in predict).  Same name over the
torch Seq2Seq here."""
from .Seq2Seq_pytorch import Seq2SeqPytorch


class LSTMSeq2Seq(Seq2SeqPytorch):
    pass
