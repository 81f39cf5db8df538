"""Image schema helpers (reference: pyzoo/zoo/pipeline/nnframes/nn_image_schema.py)."""


def with_origin_column(dataset, imageColumn="image", originColumn="origin", bigdl_type="float"):
    """Add a column holding the image's origin path."""
    out = dataset.copy()
    out[originColumn] = [im["origin"] for im in out[imageColumn]]
    return out
