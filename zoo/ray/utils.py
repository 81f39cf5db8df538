"""Helpers of the ray package (reference: pyzoo/zoo/ray/utils.py:29-80)."""
import re


def to_list(x):
    return x if isinstance(x, (list, tuple)) else [x]


def resource_to_bytes(resource_str):
    """"10g" / "512m" / "64k" / "100b" -> bytes."""
    if not resource_str:
        return resource_str
    m = re.fullmatch(r"\s*(\d+)\s*([bkmgBKMG])\s*", str(resource_str))
    if not m:
        raise Exception(f"Illegal resource string {resource_str!r}: expected a number followed by b, k, m or g")
    return int(m.group(1)) << {"b": 0, "k": 10, "m": 20, "g": 30}[m.group(2).lower()]


def is_local(sc):
    return True


def gen_shutdown_per_node(pgids, node_ips=None):
    from .process import ProcessMonitor

    def _shutdown(_iter=None):
        ProcessMonitor.clean_fn()
        return []

    return _shutdown
