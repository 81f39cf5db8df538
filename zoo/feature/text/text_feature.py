"""TextFeature (reference: pyzoo/zoo/feature/text/text_feature.py:27-120, TextFeature.scala): a dict with the keys
`text`, `label`, `uri`, `tokens`, `indexedTokens`, `sample`, `predict`."""
from __future__ import annotations


class TextFeature(dict):
    def __init__(self, text=None, label=None, uri=None, jvalue=None, bigdl_type="float"):
        super().__init__()
        if text is not None:
            self["text"] = text
        if label is not None:
            self["label"] = int(label)
        if uri is not None:
            self["uri"] = uri

    def get_text(self):
        return self.get("text")

    def get_label(self):
        return self.get("label", -1)

    def has_label(self):
        return "label" in self

    def set_label(self, label):
        self["label"] = int(label)
        return self

    def get_uri(self):
        return self.get("uri")

    def keys(self):
        return list(super().keys())

    def get_tokens(self):
        return self.get("tokens")

    def get_sample(self):
        return self.get("sample")

    def get_predict(self):
        return self.get("predict")
