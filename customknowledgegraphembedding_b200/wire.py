"""The reference's batch wire format without TensorFlow.

One TFRecord record = one *batch* (tf.train.Example with the int64 lists ``positive_sample`` [B*3],
``negative_sample`` [B*N], ``mode`` [B] and the float list ``subsampling_weight`` [B]) -- writer
/root/reference/compress_data/utils.py:35-42 + compress_data/main.py:97-125, reader
/root/reference/tensorflow_codes/run.py:40-66 (``parse_tfrecord_fn`` + ``reshape_function``).  Framing, CRC-32C and
the protobuf wire decoding run in ``libkge_b200.so`` (csrc/kge_record.cu, host code); this module maps shards into
memory and hands out torch tensors in the shapes ``reshape_function`` produces:

    positive_sample [B, 3] i64, negative_sample [B, N] i64, subsampling_weight [B, 1] f32, mode [B] i64

with ``mode`` the TF wire ints 0 = head-batch, 1 = tail-batch (compress_data/main.py:65,69).
"""
from __future__ import annotations

import ctypes as C
import glob
import os
from typing import Iterable, Iterator, List, Sequence, Tuple

import numpy as np
import torch

from . import _lib

MODE_NAMES = {0: "head-batch", 1: "tail-batch", 3: "single"}


def crc32c(data: bytes) -> int:
    buf = np.frombuffer(data, dtype=np.uint8)
    return int(_lib.load().kge_crc32c(buf.ctypes.data if buf.size else None, buf.size))


def iter_records(path: str, verify_crc: bool = True) -> Iterator[np.ndarray]:
    """Payloads of one TFRecord shard as uint8 views into a read-only memory map."""
    lib = _lib.load()
    size = os.path.getsize(path)
    if size == 0:
        return
    mm = np.memmap(path, dtype=np.uint8, mode="r")
    base = mm.ctypes.data
    off, d_off, d_len, nxt = 0, C.c_size_t(), C.c_size_t(), C.c_size_t()
    while True:
        rc = lib.kge_tfrecord_next(base, size, off, C.byref(d_off), C.byref(d_len), C.byref(nxt), int(verify_crc))
        if rc == 1:
            return
        _lib.check(rc, "kge_tfrecord_next(%s)" % path)
        yield mm[d_off.value:d_off.value + d_len.value]
        off = nxt.value


def check_id_ranges(pos, neg, nentity: int = 0, nrelation: int = 0):
    """Entity / relation ids of a parsed batch against the table sizes: ``tf.gather`` / ``index_select`` raise on an
    out-of-range id, the row-gather kernels would read (and the Adam bucket build would write) out of bounds."""
    if nentity:
        lo = min(int(pos[:, 0].min()), int(pos[:, 2].min()), int(neg.min()))
        hi = max(int(pos[:, 0].max()), int(pos[:, 2].max()), int(neg.max()))
        if lo < 0 or hi >= nentity:
            raise IndexError("entity id out of range [0, %d): batch holds ids in [%d, %d]" % (nentity, lo, hi))
    if nrelation:
        lo, hi = int(pos[:, 1].min()), int(pos[:, 1].max())
        if lo < 0 or hi >= nrelation:
            raise IndexError("relation id out of range [0, %d): batch holds ids in [%d, %d]" % (nrelation, lo, hi))


def parse_example(payload, batch_size: int, pin: bool = False, nentity: int = 0, nrelation: int = 0):
    """``parse_tfrecord_fn`` + ``reshape_function(batch_size)`` (run.py:40-66) -> four torch tensors.  With ``nentity`` /
    ``nrelation`` the ids are range-checked (IndexError, as the reference's gathers would raise)."""
    lib = _lib.load()
    buf = np.ascontiguousarray(np.frombuffer(payload, dtype=np.uint8)) if not isinstance(payload, np.ndarray) else payload
    if not buf.flags["C_CONTIGUOUS"]:
        buf = np.ascontiguousarray(buf)
    B = int(batch_size)
    mk = lambda n, dt: torch.empty(n, dtype=dt, pin_memory=pin)
    # a varint takes at least one byte: the payload length bounds the number of negatives
    pos, neg, w, mode = mk(B * 3, torch.int64), mk(max(int(buf.size), 1), torch.int64), mk(B, torch.float32), mk(B, torch.int64)
    counts = (C.c_size_t * 4)()
    _lib.check(lib.kge_record_parse_example(buf.ctypes.data, buf.size, pos.data_ptr(), pos.numel(), neg.data_ptr(),
                                            neg.numel(), w.data_ptr(), w.numel(), mode.data_ptr(), mode.numel(), counts),
               "kge_record_parse_example")
    n_pos, n_neg, n_w, n_mode = (int(c) for c in counts)
    # tf.reshape(x, [batch_size, -1]) raises when the element count is not a multiple of batch_size
    if n_pos != B * 3 or n_w != B or n_mode != B or n_neg == 0 or n_neg % B:
        raise ValueError("record does not hold a batch of %d: %d positive ids, %d negatives, %d weights, %d modes"
                         % (B, n_pos, n_neg, n_w, n_mode))
    pos, neg = pos.view(B, 3), neg[:n_neg].view(B, n_neg // B)
    if nentity or nrelation:
        check_id_ranges(pos, neg, nentity, nrelation)
    return pos, neg, w.view(B, 1), mode


def build_example(positive_sample, negative_sample, subsampling_weight, mode) -> bytes:
    """``create_example`` (compress_data/utils.py:35-42) + ``SerializeToString``: flat lists in, bytes out."""
    lib = _lib.load()
    p = np.ascontiguousarray(np.asarray(positive_sample, dtype=np.int64).reshape(-1))
    n = np.ascontiguousarray(np.asarray(negative_sample, dtype=np.int64).reshape(-1))
    w = np.ascontiguousarray(np.asarray(subsampling_weight, dtype=np.float32).reshape(-1))
    m = np.ascontiguousarray(np.asarray(mode, dtype=np.int64).reshape(-1))
    ptr = lambda a: a.ctypes.data if a.size else None
    need = lib.kge_record_build_example(ptr(p), p.size, ptr(n), n.size, ptr(w), w.size, ptr(m), m.size, None, 0)
    out = np.empty(need, dtype=np.uint8)
    got = lib.kge_record_build_example(ptr(p), p.size, ptr(n), n.size, ptr(w), w.size, ptr(m), m.size, out.ctypes.data, need)
    if got != need:
        raise RuntimeError("kge_record_build_example wrote %d of %d bytes" % (got, need))
    return out.tobytes()


def frame(payload: bytes) -> bytes:
    lib = _lib.load()
    src = np.frombuffer(payload, dtype=np.uint8)
    out = np.empty(len(payload) + 16, dtype=np.uint8)
    lib.kge_tfrecord_frame(src.ctypes.data if src.size else None, src.size, out.ctypes.data)
    return out.tobytes()


def write_tfrecords(batches: Iterable[Tuple], path: str, batch_size: int) -> int:
    """``write_file_tfrecords`` for one shard (compress_data/main.py:97-125): one record per batch; batches whose
    first dimension is not ``batch_size`` are dropped (main.py:114-116).  Returns the number of records written."""
    written = 0
    with open(path, "wb") as f:
        for positive_sample, negative_sample, subsampling_weight, mode in batches:
            p = np.asarray(positive_sample)
            if p.shape[0] != batch_size:
                continue
            m = np.asarray(mode)
            if m.ndim == 0:  # a mode name / scalar for the whole batch
                m = np.full(batch_size, int(m), dtype=np.int64)
            f.write(frame(build_example(np.hstack(p), np.hstack(np.asarray(negative_sample)),
                                        np.hstack(np.asarray(subsampling_weight)), m)))
            written += 1
    return written


class TFRecordBatchReader:
    """``TFRecordDataset(input_path).map(parse).map(reshape).repeat()`` (run.py:87-92) as a Python iterator that feeds
    ``Trainer`` / ``KGEModel.train_step``.  ``as_torch_batch=True`` yields the PyTorch-side tuple
    ``(positive [B,3], negative [B,N], weight [B], mode_name)`` instead of the TF-side one."""

    def __init__(self, paths, batch_size: int, repeat: bool = True, verify_crc: bool = True, pin: bool = False,
                 as_torch_batch: bool = False, nentity: int = 0, nrelation: int = 0):
        if isinstance(paths, str):
            paths = sorted(glob.glob(paths)) if any(ch in paths for ch in "*?[") else [paths]
        self.paths: List[str] = list(paths)
        if not self.paths:
            raise FileNotFoundError("no TFRecord shard matches")
        self.batch_size, self.repeat, self.verify_crc, self.pin = int(batch_size), repeat, verify_crc, pin
        self.as_torch_batch = as_torch_batch
        self.nentity, self.nrelation = int(nentity or 0), int(nrelation or 0)  # 0: no range check

    def __iter__(self):
        while True:
            n = 0
            for path in self.paths:
                for payload in iter_records(path, self.verify_crc):
                    pos, neg, w, mode = parse_example(payload, self.batch_size, pin=self.pin, nentity=self.nentity,
                                                        nrelation=self.nrelation)
                    n += 1
                    if self.as_torch_batch:
                        yield pos, neg, w.view(-1), MODE_NAMES[int(mode[0])]
                    else:
                        yield pos, neg, w, mode
            if not self.repeat or n == 0:
                return
