"""N2: the batch wire format (TFRecord framing + tf.train.Example) read and written without TensorFlow, against
  * golden bytes serialised by the official protobuf runtime (tests/golden/wire_golden.npz, make_wire_golden.py),
  * the RFC 3720 CRC-32C vectors,
  * the pure-Python restatement oracle/wire_oracle.py (any map order, unpacked lists, negative ids)."""
import os

import numpy as np
import pytest
import torch

from oracle import wire_oracle as wo

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "wire_golden.npz")


def test_crc32c_rfc3720_vectors():
    from customknowledgegraphembedding_b200 import wire
    for data, want in ((b"123456789", 0xE3069283), (bytes(32), 0x8A9136AA), (b"\xff" * 32, 0x62A8AB43),
                       (bytes(range(32)), 0x46DD794E), (bytes(range(31, -1, -1)), 0x113FDB5C), (b"", 0)):
        assert wire.crc32c(data) == want == wo.crc32c(data)
    blob = bytes(np.random.RandomState(0).randint(0, 256, 1001, dtype=np.uint8))  # slicing-by-8 tail handling
    for n in (1, 7, 8, 9, 63, 1001):
        assert wire.crc32c(blob[:n]) == wo.crc32c(blob[:n])


def test_golden_examples_from_the_protobuf_runtime():
    from customknowledgegraphembedding_b200 import wire
    g = np.load(GOLDEN)
    for i in range(3):
        pos, neg, w, mode = g["pos_%d" % i], g["neg_%d" % i], g["w_%d" % i], g["mode_%d" % i]
        B = pos.shape[0]
        # writer: byte-identical to protobuf's deterministic serialisation
        assert wire.build_example(pos.reshape(-1), neg.reshape(-1), w, mode) == g["example_%d" % i].tobytes()
        # reader: both the deterministic and the hash-ordered serialisation
        for key in ("example_%d", "example_any_order_%d"):
            p, n, ww, m = wire.parse_example(g[key % i].tobytes(), B)
            assert p.shape == (B, 3) and n.shape == neg.shape and ww.shape == (B, 1) and m.shape == (B,)
            assert np.array_equal(p.numpy(), pos) and np.array_equal(n.numpy(), neg)
            assert np.array_equal(ww.numpy().reshape(-1), w) and np.array_equal(m.numpy(), mode)


def test_golden_shard_and_round_trip(tmp_path):
    from customknowledgegraphembedding_b200 import wire
    g = np.load(GOLDEN)
    path = str(tmp_path / "wn18rr-0.tfrec")
    with open(path, "wb") as f:
        f.write(g["shard"].tobytes())
    recs = [bytes(r) for r in wire.iter_records(path)]
    assert recs == [g["example_%d" % i].tobytes() for i in range(3)] == wo.unframe(g["shard"].tobytes())
    # our writer produces the same shard bytes for the batches of size 4 only (others are dropped, main.py:114-116)
    batches = [(g["pos_%d" % i], g["neg_%d" % i], g["w_%d" % i], g["mode_%d" % i]) for i in range(3)]
    out = str(tmp_path / "out.tfrec")
    assert wire.write_tfrecords(batches, out, batch_size=4) == 1
    assert open(out, "rb").read() == wo.frame(g["example_0"].tobytes())
    # reader as the trainer sees it: repeat() and the PyTorch-side tuple
    it = iter(wire.TFRecordBatchReader(out, batch_size=4, repeat=True, as_torch_batch=True))
    for _ in range(3):
        p, n, w, mode = next(it)
        assert mode == "head-batch" and w.shape == (4,) and np.array_equal(n.numpy(), g["neg_0"])
    with pytest.raises(ValueError):   # tf.reshape(x, [batch_size, -1]) would fail
        list(wire.TFRecordBatchReader(path, batch_size=4, repeat=False))


def test_reader_accepts_every_legal_encoding_and_rejects_damage(tmp_path):
    from customknowledgegraphembedding_b200 import wire
    pos, neg, w, mode = [3, 0, 9, 5, 1, 2], [-1, 2 ** 40, 7, 0, 1, 300], [0.25, 0.5], [1, 1]
    for order in (("positive_sample", "mode", "subsampling_weight", "negative_sample"),
                  ("mode", "negative_sample", "positive_sample", "subsampling_weight")):
        for packed in (True, False):
            buf = wo.build_example(pos, neg, w, mode, order=order, packed=packed)
            p, n, ww, m = wire.parse_example(buf, 2)
            assert p.reshape(-1).tolist() == pos and n.reshape(-1).tolist() == neg
            assert ww.reshape(-1).tolist() == w and m.tolist() == mode
            assert wo.parse_example(buf) == {"positive_sample": pos, "negative_sample": neg, "subsampling_weight": w, "mode": mode}
    good = wo.frame(wo.build_example(pos, neg, w, mode))
    path = str(tmp_path / "x.tfrec")
    for damage in ("data", "length", "truncated"):
        blob = bytearray(good)
        if damage == "data":
            blob[20] ^= 1
        elif damage == "length":
            blob[0] ^= 1
        else:
            blob = blob[:-3]
        with open(path, "wb") as f:
            f.write(bytes(blob))
        with pytest.raises(ValueError):
            list(wire.iter_records(path))
    with open(path, "wb") as f:
        f.write(b"")
    assert list(wire.iter_records(path)) == []
    with pytest.raises(ValueError):
        wire.parse_example(b"\x0a\x05\x0a\x03", 2)   # truncated message


def test_full_size_record_round_trip():
    """BASELINE configs[1] batch: 1024 x 256 negatives, ids up to 14541 -- 2-byte varints, ~0.5 MB per record."""
    from customknowledgegraphembedding_b200 import wire
    rng = np.random.RandomState(1)
    B, N, E = 1024, 256, 14541
    pos = np.stack([rng.randint(0, E, B), rng.randint(0, 237, B), rng.randint(0, E, B)], 1)
    neg = rng.randint(0, E, (B, N))
    w = np.sqrt(1.0 / rng.randint(8, 200, B)).astype(np.float32)
    buf = wire.build_example(pos.reshape(-1), neg.reshape(-1), w, np.ones(B, dtype=np.int64))
    p, n, ww, m = wire.parse_example(buf, B)
    assert np.array_equal(p.numpy(), pos) and np.array_equal(n.numpy(), neg) and np.array_equal(ww.numpy().reshape(-1), w)
    assert int(m.sum()) == B
