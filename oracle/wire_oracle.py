"""Test infrastructure: pure-Python restatement of the batch wire format (TFRecord framing + tf.train.Example wire
encoding) used to check customknowledgegraphembedding_b200/wire.py + csrc/kge_record.cu.

Follows /root/reference/compress_data/utils.py:35-42 (create_example), compress_data/main.py:113-125 (one record per
batch) and tensorflow_codes/run.py:40-66 (parse + reshape).  TensorFlow is not installable here; the formats are the
public ones: TFRecord (u64 length, masked CRC-32C of the length, data, masked CRC-32C of the data) and protobuf wire
format of tensorflow/core/example/{example,feature}.proto.  Pinned by: RFC 3720 CRC-32C vectors, and messages
serialised by the official ``google.protobuf`` runtime from the same schema (tests/test_wire.py).
Only tests/ may import this module."""
import struct


def crc32c(data: bytes) -> int:
    c = 0xFFFFFFFF
    for b in data:
        c ^= b
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
    return c ^ 0xFFFFFFFF


def masked_crc(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def frame(payload: bytes) -> bytes:
    head = struct.pack("<Q", len(payload))
    return head + struct.pack("<I", masked_crc(head)) + payload + struct.pack("<I", masked_crc(payload))


def unframe(blob: bytes):
    out, off = [], 0
    while off < len(blob):
        (n,) = struct.unpack_from("<Q", blob, off)
        assert struct.unpack_from("<I", blob, off + 8)[0] == masked_crc(blob[off:off + 8])
        data = blob[off + 12:off + 12 + n]
        assert struct.unpack_from("<I", blob, off + 12 + n)[0] == masked_crc(data)
        out.append(data)
        off += 16 + n
    return out


def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _len_field(field: int, payload: bytes) -> bytes:
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def build_example(positive, negative, weight, mode, order=("mode", "negative_sample", "positive_sample", "subsampling_weight"),
                  packed=True) -> bytes:
    def int_list(vals):
        if packed:
            return _len_field(1, b"".join(_varint(int(v)) for v in vals)) if len(vals) else b""
        return b"".join(_varint(1 << 3) + _varint(int(v)) for v in vals)

    def float_list(vals):
        if packed:
            return _len_field(1, struct.pack("<%df" % len(vals), *vals)) if len(vals) else b""
        return b"".join(_varint((1 << 3) | 5) + struct.pack("<f", v) for v in vals)
    feats = {"positive_sample": _len_field(3, int_list(positive)), "negative_sample": _len_field(3, int_list(negative)),
             "subsampling_weight": _len_field(2, float_list(weight)), "mode": _len_field(3, int_list(mode))}
    body = b"".join(_len_field(1, _len_field(1, k.encode()) + _len_field(2, feats[k])) for k in order)
    return _len_field(1, body)


def _read_varint(buf, off):
    v, shift = 0, 0
    while True:
        b = buf[off]
        off += 1
        v |= (b & 0x7F) << shift
        shift += 7
        if not b & 0x80:
            return v & ((1 << 64) - 1), off


def _fields(buf):
    off = 0
    while off < len(buf):
        tag, off = _read_varint(buf, off)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            v, off = _read_varint(buf, off)
        elif wt == 2:
            n, off = _read_varint(buf, off)
            v, off = buf[off:off + n], off + n
        elif wt == 5:
            v, off = buf[off:off + 4], off + 4
        elif wt == 1:
            v, off = buf[off:off + 8], off + 8
        else:
            raise ValueError("wire type %d" % wt)
        yield field, wt, v


def parse_example(buf: bytes):
    """-> dict name -> list of python ints / floats."""
    out = {}
    for f, wt, features in _fields(buf):
        if f != 1 or wt != 2:
            continue
        for f1, wt1, entry in _fields(features):
            if f1 != 1 or wt1 != 2:
                continue
            key, feature = None, b""
            for f2, wt2, v in _fields(entry):
                if f2 == 1:
                    key = bytes(v).decode()
                elif f2 == 2:
                    feature = v
            vals = []
            for kind, wtk, lst in _fields(feature):
                for f3, wt3, v in _fields(lst):
                    if f3 != 1:
                        continue
                    if kind == 3:
                        if wt3 == 2:
                            off = 0
                            while off < len(v):
                                x, off = _read_varint(v, off)
                                vals.append(x - (1 << 64) if x >= 1 << 63 else x)
                        else:
                            vals.append(v - (1 << 64) if v >= 1 << 63 else v)
                    elif kind == 2:
                        if wt3 == 2:
                            vals.extend(struct.unpack("<%df" % (len(v) // 4), bytes(v)))
                        else:
                            vals.append(struct.unpack("<f", bytes(v))[0])
            out[key] = vals
    return out
