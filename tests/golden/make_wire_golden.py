"""Generates tests/golden/wire_golden.npz: tf.train.Example messages serialised by the OFFICIAL protobuf runtime
(google.protobuf, from a descriptor built to tensorflow/core/example/{example,feature}.proto) for the batch record of
/root/reference/compress_data/utils.py:35-42, framed as TFRecord with the pure-Python oracle (CRC-32C checked against
RFC 3720).  TensorFlow itself cannot be installed here; these bytes are what tf.train.Example.SerializeToString /
tf.io.TFRecordWriter emit for the same schema.  Run: python tests/golden/make_wire_golden.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import wire_oracle as wo  # noqa: E402


def example_class():
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fd = descriptor_pb2.FileDescriptorProto(name="kge_example.proto", package="tensorflow", syntax="proto3")
    T = descriptor_pb2.FieldDescriptorProto

    def msg(name):
        m = fd.message_type.add()
        m.name = name
        return m

    def field(m, name, number, ftype, label=T.LABEL_OPTIONAL, type_name=None, packed=None, oneof=None):
        f = m.field.add(name=name, number=number, type=ftype, label=label)
        if type_name:
            f.type_name = type_name
        if packed is not None:
            f.options.packed = packed
        if oneof is not None:
            f.oneof_index = oneof
        return f
    field(msg("BytesList"), "value", 1, T.TYPE_BYTES, T.LABEL_REPEATED)
    field(msg("FloatList"), "value", 1, T.TYPE_FLOAT, T.LABEL_REPEATED, packed=True)
    field(msg("Int64List"), "value", 1, T.TYPE_INT64, T.LABEL_REPEATED, packed=True)
    feature = msg("Feature")
    feature.oneof_decl.add(name="kind")
    field(feature, "bytes_list", 1, T.TYPE_MESSAGE, type_name=".tensorflow.BytesList", oneof=0)
    field(feature, "float_list", 2, T.TYPE_MESSAGE, type_name=".tensorflow.FloatList", oneof=0)
    field(feature, "int64_list", 3, T.TYPE_MESSAGE, type_name=".tensorflow.Int64List", oneof=0)
    features = msg("Features")
    entry = features.nested_type.add(name="FeatureEntry")
    entry.options.map_entry = True
    field(entry, "key", 1, T.TYPE_STRING)
    field(entry, "value", 2, T.TYPE_MESSAGE, type_name=".tensorflow.Feature")
    field(features, "feature", 1, T.TYPE_MESSAGE, T.LABEL_REPEATED, type_name=".tensorflow.Features.FeatureEntry")
    field(msg("Example"), "features", 1, T.TYPE_MESSAGE, type_name=".tensorflow.Features")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("tensorflow.Example"))


def make_example(cls, positive, negative, weight, mode):
    ex = cls()
    f = ex.features.feature
    f["positive_sample"].int64_list.value.extend(int(x) for x in positive)
    f["negative_sample"].int64_list.value.extend(int(x) for x in negative)
    f["subsampling_weight"].float_list.value.extend(float(x) for x in weight)
    f["mode"].int64_list.value.extend(int(x) for x in mode)
    return ex


def batches():
    rng = np.random.RandomState(0)
    out = []
    for B, N, E in ((4, 6, 40943), (8, 3, 10_000_000), (2, 5, 100)):
        pos = np.stack([rng.randint(0, E, B), rng.randint(0, 11, B), rng.randint(0, E, B)], 1).astype(np.int64)
        neg = rng.randint(0, E, (B, N)).astype(np.int64)
        w = np.sqrt(1.0 / rng.randint(8, 200, B)).astype(np.float32)
        mode = np.full(B, len(out) % 2, dtype=np.int64)
        out.append((pos, neg, w, mode))
    return out


if __name__ == "__main__":
    cls = example_class()
    blobs, shard = {}, b""
    for i, (pos, neg, w, mode) in enumerate(batches()):
        ex = make_example(cls, pos.reshape(-1), neg.reshape(-1), w, mode)
        det = ex.SerializeToString(deterministic=True)
        blobs["example_%d" % i] = np.frombuffer(det, dtype=np.uint8)
        blobs["example_any_order_%d" % i] = np.frombuffer(ex.SerializeToString(), dtype=np.uint8)
        for k, v in (("pos", pos), ("neg", neg), ("w", w), ("mode", mode)):
            blobs["%s_%d" % (k, i)] = v
        shard += wo.frame(det)
    blobs["shard"] = np.frombuffer(shard, dtype=np.uint8)
    here = os.path.dirname(os.path.abspath(__file__))
    np.savez(os.path.join(here, "wire_golden.npz"), **blobs)
    print("wrote wire_golden.npz", {k: v.shape for k, v in blobs.items() if k.startswith("example")})
