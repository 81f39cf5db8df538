"""Keras-style dataset loaders (reference: pyzoo/zoo/pipeline/api/keras/datasets/{mnist,imdb,reuters,boston_housing}.py).
The reference downloads the files; this box has no network, so each loader reads the standard files from a local
directory when they are there and otherwise returns synthetic data of the same shapes and dtypes (`synthetic=True`
is printed once)."""
