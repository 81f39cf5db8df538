"""Data layer: Sample, Preprocessing chain, Relations and FeatureSet.

Reference: pyzoo/zoo/feature/common.py:28-379 and zoo/src/main/scala/.../feature/FeatureSet.scala
(53-709), feature/common/*.scala.  The reference caches one array per Spark partition and
serves an *infinite looped iterator* with a random offset for training and a finite one for
evaluation (CachedDistributedFeatureSet, FeatureSet.scala:230-330), optionally keeping the
cache in persistent memory (PMEM/DIRECT) or sampling 1/n of a disk-resident set per pass
(DiskFeatureSet, :332-431).  On one multi-GPU box:

=================  =====================================================================
memory_type        where the cache lives
=================  =====================================================================
``DRAM``           pinned host memory (page-locked, so H2D copies are asynchronous DMA)
``PMEM``/``DIRECT``/``DEVICE``  resident in the GPU's HBM (180 GB per B200): zero copies per step
``DISK_AND_DRAM(n)``  ``numpy.memmap`` on disk; 1/n of the samples is staged to pinned memory per slice
=================  =====================================================================

Each rank holds only its shard (`FeatureSet.shard(rank, world)`), the analogue of
``coalesce(nodeNumber)`` (FeatureSet.scala:677-678).
"""
from __future__ import annotations

import csv
import math
import os
import re
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

__all__ = [
    "Sample", "Preprocessing", "ChainedPreprocessing", "ScalarToTensor", "SeqToTensor", "SeqToMultipleTensors",
    "ArrayToTensor", "MLlibVectorToTensor", "FeatureLabelPreprocessing", "TensorToSample", "ToTuple",
    "SampleToMiniBatch", "MTSampleToMiniBatch", "FeatureToTupleAdapter", "BigDLAdapter", "Relation", "Relations",
    "FeatureSet", "MiniBatch", "MemoryType",
]


class Sample:
    """features: ndarray or list of ndarrays; label(s) likewise (BigDL `Sample.from_ndarray`)."""

    __slots__ = ("features", "labels")

    def __init__(self, features, labels=None):
        self.features = [np.asarray(f) for f in features] if isinstance(features, (list, tuple)) else [np.asarray(features)]
        if labels is None:
            self.labels = []
        else:
            self.labels = [np.asarray(l) for l in labels] if isinstance(labels, (list, tuple)) else [np.asarray(labels)]

    @classmethod
    def from_ndarray(cls, features, labels=None, bigdl_type="float"):
        return cls(features, labels)

    @property
    def feature(self):
        return self.features[0]

    @property
    def label(self):
        return self.labels[0] if self.labels else None

    def __repr__(self):
        return f"Sample(features={[f.shape for f in self.features]}, labels={[l.shape for l in self.labels]})"


class MiniBatch:
    """(inputs, targets) of one step; inputs/targets are tensors or tuples of tensors."""

    __slots__ = ("inputs", "targets", "size", "_h2d_done")

    def __init__(self, inputs, targets, size):
        self.inputs, self.targets, self.size = inputs, targets, size
        self._h2d_done = None     # CUDA event recorded by the prefetcher once the batch has been copied to the device

    def to(self, device, non_blocking=True):
        mv = lambda t: None if t is None else (tuple(x.to(device, non_blocking=non_blocking) for x in t)
                                               if isinstance(t, (tuple, list)) else t.to(device, non_blocking=non_blocking))
        return MiniBatch(mv(self.inputs), mv(self.targets), self.size)


# ------------------------------------------------------------- preprocessing --
class Preprocessing:
    """A -> B transformer; chain with `ChainedPreprocessing([...])` or `a >> b`."""

    def __call__(self, x):
        return self.apply(x)

    def apply(self, x):
        return x

    def __rshift__(self, other):
        return ChainedPreprocessing([self, other])

    def map(self, iterable):
        for x in iterable:
            yield self.apply(x)


class ChainedPreprocessing(Preprocessing):
    def __init__(self, transformers):
        flat = []
        for t in transformers:
            if isinstance(t, ChainedPreprocessing):
                flat.extend(t.transformers)
            else:
                flat.append(t)
        for t in flat:
            if not callable(t):
                raise TypeError(f"{t!r} is not a Preprocessing")
        self.transformers = flat

    def apply(self, x):
        for t in self.transformers:
            x = t(x)
        return x


class ScalarToTensor(Preprocessing):
    def apply(self, x):
        return np.asarray([x], dtype=np.float32)


class SeqToTensor(Preprocessing):
    def __init__(self, size=None):
        self.size = list(size) if size else None

    def apply(self, x):
        a = np.asarray(x, dtype=np.float32)
        return a.reshape(self.size) if self.size else a


class SeqToMultipleTensors(Preprocessing):
    """Split one flat sequence into several tensors of the given sizes."""

    def __init__(self, size):
        self.sizes = [list(s) for s in size]

    def apply(self, x):
        a, out, off = np.asarray(x, dtype=np.float32).ravel(), [], 0
        for s in self.sizes:
            n = int(np.prod(s))
            out.append(a[off:off + n].reshape(s))
            off += n
        return out


class ArrayToTensor(SeqToTensor):
    pass


class MLlibVectorToTensor(SeqToTensor):
    def apply(self, x):
        return super().apply(x.toArray() if hasattr(x, "toArray") else x)


class TensorToSample(Preprocessing):
    def apply(self, x):
        return x if isinstance(x, Sample) else Sample(x)


class MultiTensorsToSample(Preprocessing):
    """A list of tensors -> one Sample whose feature is all of them, no label (reference:
    feature/common/MultiTensorsToSample.scala:27-33; used behind `SeqToMultipleTensors` for multi-input models)."""

    def apply(self, x):
        if isinstance(x, Sample):
            return x
        return Sample(list(x) if isinstance(x, (tuple, list)) else [x])


class FeatureLabelPreprocessing(Preprocessing):
    """(feature, label) -> Sample using one transformer per side (common.py:188-200)."""

    def __init__(self, feature_transformer, label_transformer):
        self.ft, self.lt = feature_transformer, label_transformer

    def apply(self, x):
        f, l = x
        return Sample(self.ft(f), None if l is None else self.lt(l))


class ToTuple(Preprocessing):
    def apply(self, x):
        return (x, None)


class FeatureToTupleAdapter(Preprocessing):
    def __init__(self, sample_transformer):
        self.t = sample_transformer

    def apply(self, x):
        return self.t(x[0])


class BigDLAdapter(Preprocessing):
    def __init__(self, bigdl_transformer):
        self.t = bigdl_transformer

    def apply(self, x):
        return self.t(x)


def _stack(arrs):
    a0 = arrs[0]
    if isinstance(a0, torch.Tensor):
        return torch.stack(arrs)
    return torch.from_numpy(np.stack([np.asarray(a) for a in arrs]))


def collate_samples(samples):
    nf, nl = len(samples[0].features), len(samples[0].labels)
    feats = tuple(_stack([s.features[i] for s in samples]) for i in range(nf))
    labs = tuple(_stack([s.labels[i] for s in samples]) for i in range(nl))
    return MiniBatch(feats[0] if nf == 1 else feats, None if nl == 0 else (labs[0] if nl == 1 else labs), len(samples))


class SampleToMiniBatch(Preprocessing):
    """Group an iterator of Samples into MiniBatches."""

    def __init__(self, batch_size, drop_last=False):
        self.batch_size, self.drop_last = int(batch_size), drop_last

    def map(self, iterable):
        buf = []
        for s in iterable:
            buf.append(s)
            if len(buf) == self.batch_size:
                yield collate_samples(buf)
                buf = []
        if buf and not self.drop_last:
            yield collate_samples(buf)

    def apply(self, x):
        return collate_samples(list(x))


class MTSampleToMiniBatch(SampleToMiniBatch):
    """Multi-threaded variant: per-sample `transformer` runs on a thread pool while batching
    (feature/common/MTSampleToMiniBatch.scala)."""

    def __init__(self, batch_size, transformer=None, num_threads=None, drop_last=False):
        super().__init__(batch_size, drop_last)
        self.transformer = transformer
        self.pool = ThreadPoolExecutor(num_threads or min(16, os.cpu_count() or 4))

    def map(self, iterable):
        it = iter(iterable)
        while True:
            raw = []
            for x in it:
                raw.append(x)
                if len(raw) == self.batch_size:
                    break
            if not raw or (len(raw) < self.batch_size and self.drop_last):
                return
            samples = list(self.pool.map(self.transformer, raw)) if self.transformer else raw
            yield collate_samples(samples)
            if len(raw) < self.batch_size:
                return


# ------------------------------------------------------------------ relations --
class Relation:
    """(id1, id2, label) text-pair relation (common.py:28-60)."""

    def __init__(self, id1, id2, label):
        self.id1, self.id2, self.label = str(id1), str(id2), int(label)

    def to_tuple(self):
        return self.id1, self.id2, self.label

    def __repr__(self):
        return f"Relation [id1: {self.id1}, id2: {self.id2}, label: {self.label}]"

    def __eq__(self, o):
        return isinstance(o, Relation) and self.to_tuple() == o.to_tuple()

    def __hash__(self):
        return hash(self.to_tuple())


class Relations:
    @staticmethod
    def read(path, sc=None, min_partitions=1):
        """csv/txt file (or directory of them) with lines `id1,id2,label`; header lines are skipped."""
        files = [os.path.join(path, f) for f in sorted(os.listdir(path))] if os.path.isdir(path) else [path]
        out = []
        for fp in files:
            with open(fp, newline="") as f:
                for row in csv.reader(f):
                    if len(row) < 3:
                        continue
                    try:
                        out.append(Relation(row[0], row[1], int(row[2])))
                    except ValueError:
                        continue
        return out

    @staticmethod
    def read_parquet(path, sc=None):
        import pyarrow.parquet as pq

        t = pq.read_table(path).to_pydict()
        return [Relation(a, b, c) for a, b, c in zip(t["id1"], t["id2"], t["label"])]

    @staticmethod
    def generate_relation_pairs(relations):
        """For every id1, pair each positive (label > 0) id2 with each negative id2."""
        by = {}
        for r in relations:
            by.setdefault(r.id1, ([], []))[0 if r.label > 0 else 1].append(r.id2)
        return [(i1, p, n) for i1, (pos, neg) in by.items() for p in pos for n in neg]


# ------------------------------------------------------------------ FeatureSet --
class MemoryType:
    DRAM, PMEM, DIRECT, DEVICE = "DRAM", "PMEM", "DIRECT", "DEVICE"

    @staticmethod
    def parse(mt):
        """'DRAM' | 'PMEM' | 'DIRECT' | 'DEVICE' | 'DISK_AND_DRAM(n)' | int n  ->  (kind, n_slice)"""
        if isinstance(mt, int):
            return "DISK_AND_DRAM", max(1, mt)
        s = str(mt).upper().strip()
        m = re.match(r"DISK_AND_DRAM\(?\s*(\d+)\s*\)?", s)
        if m:
            return "DISK_AND_DRAM", max(1, int(m.group(1)))
        if s in ("PMEM", "DIRECT", "DEVICE"):
            return "DEVICE", 1
        if s == "DRAM":
            return "DRAM", 1
        raise ValueError(f"unknown memory_type {mt!r}")


def _as_tensor_tuple(x):
    if x is None:
        return None
    if isinstance(x, (list, tuple)):
        return tuple(_to_tensor(a) for a in x)
    return (_to_tensor(x),)


def _to_tensor(a):
    if isinstance(a, torch.Tensor):
        return a
    a = np.asarray(a)
    if a.dtype == np.float64:
        a = a.astype(np.float32)
    return torch.from_numpy(np.ascontiguousarray(a))


class FeatureSet:
    """Cached, shardable, optionally device-resident data set.

    Two storage modes: *tensor mode* (columns of equally long tensors; batches are index
    gathers -- the fast path) and *sample mode* (a list of arbitrary records plus a
    Preprocessing chain ending in Samples; batches are collated on a thread pool).
    """

    def __init__(self, features=None, labels=None, records=None, memory_type="DRAM", sequential_order=False,
                 shuffle=True, transformer=None, device=None, seed=1234):
        self.kind, self.num_slice = MemoryType.parse(memory_type)
        self.memory_type = memory_type
        self.sequential_order, self.do_shuffle = sequential_order, shuffle
        self.transformer = transformer
        self.features = _as_tensor_tuple(features)
        self.labels = _as_tensor_tuple(labels)
        self.single_feature = features is not None and not isinstance(features, (list, tuple))
        self.single_label = labels is not None and not isinstance(labels, (list, tuple))
        self.records = list(records) if records is not None else None
        self.device = torch.device(device) if device is not None else None
        self._rng = np.random.RandomState(seed)
        self._perm = None
        self._cached = False
        self.current_slice = 0
        self._lock = threading.Lock()
        if self.features is not None:
            n = self.features[0].shape[0]
            for t in self.features + (self.labels or ()):
                if t.shape[0] != n:
                    raise ValueError("all feature/label columns must have the same first dimension")

    # ---- constructors (names of pyzoo/zoo/feature/common.py:238-379) ---------------
    @classmethod
    def numpy(cls, x, y=None, **kw):
        return cls(features=x, labels=y, **kw)

    @classmethod
    def sample_rdd(cls, rdd, memory_type="DRAM", sequential_order=False, shuffle=True, **kw):
        """`rdd`: iterable of Sample.  Samples with uniform shapes are packed into tensor mode."""
        samples = list(rdd)
        if samples and isinstance(samples[0], Sample):
            try:
                nf, nl = len(samples[0].features), len(samples[0].labels)
                feats = [np.stack([s.features[i] for s in samples]) for i in range(nf)]
                labs = [np.stack([s.labels[i] for s in samples]) for i in range(nl)]
                return cls(features=feats[0] if nf == 1 else feats, labels=None if nl == 0 else (labs[0] if nl == 1 else labs),
                           memory_type=memory_type, sequential_order=sequential_order, shuffle=shuffle, **kw)
            except ValueError:
                pass
        return cls(records=samples, memory_type=memory_type, sequential_order=sequential_order, shuffle=shuffle, **kw)

    @classmethod
    def rdd(cls, rdd, memory_type="DRAM", sequential_order=False, shuffle=True, **kw):
        """Arbitrary records (use `.transform(...)` to turn them into Samples)."""
        return cls(records=list(rdd), memory_type=memory_type, sequential_order=sequential_order, shuffle=shuffle, **kw)

    @classmethod
    def image_set(cls, imageset, memory_type="DRAM", **kw):
        return cls(records=list(imageset.get_image_features()), memory_type=memory_type,
                   transformer=getattr(imageset, "transformer", None), **kw)

    image_frame = image_set

    @classmethod
    def pytorch_dataloader(cls, dataloader, **kw):
        """Wrap a torch DataLoader/Dataset: batches come from the loader itself
        (reference PythonLoaderFeatureSet, FeatureSet.scala:433-555, runs the loader inside jep)."""
        fs = cls(records=[], **kw)
        fs._loader = dataloader
        return fs

    @classmethod
    def xshards(cls, shards, feature_cols=None, label_cols=None, **kw):
        data = shards.collect_local() if hasattr(shards, "collect_local") else shards.collect()
        xs, ys = [], []
        for part in data:
            if isinstance(part, dict):
                xs.append(part["x"]); ys.append(part.get("y"))
            else:  # pandas DataFrame
                xs.append(part[feature_cols].to_numpy()); ys.append(part[label_cols].to_numpy() if label_cols else None)
        cat = lambda parts: (tuple(np.concatenate([p[i] for p in parts]) for i in range(len(parts[0])))
                             if isinstance(parts[0], (tuple, list)) else np.concatenate(parts))
        y = None if ys[0] is None else cat(ys)
        x = cat(xs)
        return cls(features=list(x) if isinstance(x, tuple) else x, labels=list(y) if isinstance(y, tuple) else y, **kw)

    @classmethod
    def tf_dataset(cls, func, total_size=None, bigdl_type="float"):
        """FeatureSet over a tf.data-style source (reference: FeatureSet.tf_dataset(func, total_size), common.py:330-345 -- `func` builds
        the dataset on each executor).  `func` is a callable returning an iterable of elements (or the iterable itself); elements are
        arrays or (features, labels) pairs; at most `total_size` of them are read."""
        import itertools

        src = func() if callable(func) else func
        it = src.as_numpy_iterator() if hasattr(src, "as_numpy_iterator") else iter(src)
        elems = list(itertools.islice(it, total_size) if total_size is not None else it)
        if not elems:
            raise ValueError("FeatureSet.tf_dataset: the dataset is empty")
        if isinstance(elems[0], (tuple, list)) and len(elems[0]) == 2:
            return cls.numpy(np.stack([np.asarray(e[0]) for e in elems]), np.stack([np.asarray(e[1]) for e in elems]))
        return cls.numpy(np.stack([np.asarray(e) for e in elems]))

    # ---- basic ops ---------------------------------------------------------------
    def size(self):
        if self.features is not None:
            return int(self.features[0].shape[0])
        if getattr(self, "_loader", None) is not None:
            return len(self._loader.dataset) if hasattr(self._loader, "dataset") else len(self._loader)
        return len(self.records)

    __len__ = size

    def transform(self, transformer):
        """Lazy per-record transformer (chains with an existing one)."""
        fs = self._clone()
        fs.transformer = transformer if fs.transformer is None else ChainedPreprocessing([fs.transformer, transformer])
        return fs

    __rshift__ = transform

    def _clone(self):
        fs = FeatureSet.__new__(FeatureSet)
        fs.__dict__.update(self.__dict__)
        fs._lock = threading.Lock()
        return fs

    def to_dataset(self):
        return self

    def shard(self, rank, world):
        """Keep the `rank`-th of `world` interleaved shards (equal sizes: tail dropped)."""
        if world <= 1:
            return self
        fs = self._clone()
        n = self.size() // world * world
        if self.features is not None:
            fs.features = tuple(t[rank:n:world] for t in self.features)
            fs.labels = None if self.labels is None else tuple(t[rank:n:world] for t in self.labels)
        elif self.records is not None:
            fs.records = self.records[rank:n:world]
        fs._perm, fs._cached = None, False
        return fs

    def cache(self, device=None):
        """Materialise the cache in its memory tier."""
        if self._cached or self.features is None:
            return self
        if self.kind == "DEVICE":
            dev = torch.device(device) if device is not None else (self.device or (torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")))
            self.features = tuple(t.to(dev) for t in self.features)
            self.labels = None if self.labels is None else tuple(t.to(dev) for t in self.labels)
            self.device = dev
        elif self.kind == "DRAM" and torch.cuda.is_available():
            pin = lambda t: t if t.is_pinned() or t.is_cuda else t.contiguous().pin_memory()
            self.features = tuple(pin(t) for t in self.features)
            self.labels = None if self.labels is None else tuple(pin(t) for t in self.labels)
        self._cached = True
        return self

    def unpersist(self):
        self._cached = False
        return self

    def shuffle(self):
        n = self._slice_len()
        self._perm = self._rng.permutation(n)
        return self

    def _slice_len(self):
        n = self.size()
        return n if self.num_slice <= 1 else int(math.ceil(n / self.num_slice))

    def num_of_slice(self):
        return self.num_slice

    numOfSlice = num_of_slice

    # ---- iteration ---------------------------------------------------------------
    def _gather(self, idx):
        sel = lambda t: t[idx] if not torch.is_tensor(idx) else t.index_select(0, idx.to(t.device))
        f = tuple(sel(t) for t in self.features)
        l = None if self.labels is None else tuple(sel(t) for t in self.labels)
        return MiniBatch(f[0] if self.single_feature else f, None if l is None else (l[0] if self.single_label else l), len(idx))

    def _window(self, lo, n):
        """Contiguous batch: zero-copy views (of pinned memory for the DRAM tier, so the H2D copy is one async DMA)."""
        f = tuple(t.narrow(0, lo, n) for t in self.features)
        l = None if self.labels is None else tuple(t.narrow(0, lo, n) for t in self.labels)
        return MiniBatch(f[0] if self.single_feature else f, None if l is None else (l[0] if self.single_label else l), n)

    def _gather_pinned(self, idx):
        """Shuffled batch from host memory: gathered into a small ring of pinned staging buffers so the H2D copy stays
        asynchronous.  A slot is reused only after the copy that read it has finished (`_h2d_done`, set by the prefetcher)."""
        ring = self.__dict__.setdefault("_stage_ring", {"slots": [None] * 4, "i": 0})
        k = ring["i"]
        ring["i"] = (k + 1) % len(ring["slots"])
        slot = ring["slots"][k]
        n = len(idx)
        tensors = list(self.features) + (list(self.labels) if self.labels is not None else [])
        if slot is None or slot["n"] < n:
            slot = ring["slots"][k] = {"n": n, "bufs": [torch.empty((n,) + tuple(t.shape[1:]), dtype=t.dtype).pin_memory() for t in tensors],
                                       "batch": None}
        old = slot["batch"]
        ev = getattr(old, "_h2d_done", None) if old is not None else None
        if ev is not None:
            ev.synchronize()
        outs = []
        for t, buf in zip(tensors, slot["bufs"]):
            o = buf.narrow(0, 0, n)
            torch.index_select(t, 0, idx, out=o)
            outs.append(o)
        nf = len(self.features)
        f, l = tuple(outs[:nf]), (None if self.labels is None else tuple(outs[nf:]))
        b = MiniBatch(f[0] if self.single_feature else f, None if l is None else (l[0] if self.single_label else l), n)
        slot["batch"] = b
        return b

    def _slice_offset(self):
        if self.num_slice <= 1:
            return 0, self.size()
        per = self._slice_len()
        s = self.current_slice % self.num_slice
        return s * per, min(self.size(), (s + 1) * per)

    def batches(self, batch_size, train=False, drop_last=None, epochs=1):
        """Finite iterator over `epochs` passes (a 'pass' is one slice for DISK_AND_DRAM sets)."""
        drop_last = train if drop_last is None else drop_last
        loader = getattr(self, "_loader", None)
        if loader is not None:
            for _ in range(epochs):
                for b in loader:
                    x, y = (b[0], b[1]) if isinstance(b, (list, tuple)) and len(b) >= 2 else (b, None)
                    yield MiniBatch(x, y, (x[0] if isinstance(x, (list, tuple)) else x).shape[0])
            return
        if self.features is not None:
            self.cache()
            for _ in range(epochs):
                lo, hi = self._slice_offset()
                n = hi - lo
                stop = n - (n % batch_size) if drop_last else n
                dev = self.features[0].device
                if not (train and self.do_shuffle and not self.sequential_order):
                    for s in range(0, stop, batch_size):
                        yield self._window(lo + s, min(batch_size, n - s))
                else:
                    perm_t = torch.from_numpy(self._rng.permutation(n) + lo).to(dev)
                    host_pinned = dev.type == "cpu" and self.features[0].is_pinned()
                    for s in range(0, stop, batch_size):
                        idx = perm_t[s:s + batch_size]
                        yield self._gather_pinned(idx) if host_pinned else self._gather(idx)
                if self.num_slice > 1:
                    self.current_slice += 1
            return
        # sample mode
        for _ in range(epochs):
            recs = self.records
            order = self._rng.permutation(len(recs)) if (train and self.do_shuffle and not self.sequential_order) else range(len(recs))
            mt = MTSampleToMiniBatch(batch_size, self._record_to_sample, drop_last=drop_last)
            yield from mt.map(recs[i] for i in order)

    def _record_to_sample(self, r):
        if self.transformer is not None:
            r = self.transformer(r)
        if isinstance(r, Sample):
            return r
        if isinstance(r, (tuple, list)) and len(r) == 2:
            return Sample(r[0], r[1])
        return Sample(r)

    def data(self, train=True, batch_size=None):
        """Reference semantics (FeatureSet.scala:248-296): `train=True` is an endless looped
        iterator (reshuffled every pass), `train=False` a single finite pass."""
        if batch_size is None:
            it = self._iter_records(train)
            return it
        if not train:
            return self.batches(batch_size, train=False)

        def forever():
            while True:
                yield from self.batches(batch_size, train=True)

        return forever()

    def _iter_records(self, train):
        n = self.size()
        while True:
            order = self._rng.permutation(n) if (train and self.do_shuffle) else range(n)
            for i in order:
                if self.features is not None:
                    b = self._gather(np.asarray([i]))
                    yield b
                else:
                    yield self._record_to_sample(self.records[i])
            if not train:
                return
