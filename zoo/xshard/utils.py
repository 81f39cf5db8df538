"""Helpers of the xshard package (reference: pyzoo/zoo/xshard/utils.py)."""


def chunk(lst, n):
    """Split `lst` into `n` contiguous chunks of (almost) equal size."""
    from .shard import _chunks

    return iter(_chunks(list(lst), n))


def flatten(list_of_list):
    return [item for sub in list_of_list for item in sub]
