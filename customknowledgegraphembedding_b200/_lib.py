"""ctypes binding of ``libkge_b200.so`` (C-ABI declared in ``include/kge_b200.h``).

There is no fallback: if the library is missing or a call fails, this raises.  The library is
built in-tree by ``python -m customknowledgegraphembedding_b200.build`` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkge_b200.so")

ABI_VERSION = 13

# kge_model_id / kge_mode / kge_ws_kind
MODEL_IDS = {"TransE": 0, "DistMult": 1, "ComplEx": 2, "RotatE": 3, "pRotatE": 4, "InterHT": 5}
MODE_IDS = {"single": 0, "head-batch": 1, "tail-batch": 2}
WS_BACKWARD, WS_LOSS, WS_EVAL, WS_TRAIN_STEP, WS_TRAIN_STEP_HOST, WS_TRAIN_APPLY, WS_TRAIN_SWEEP = range(7)

KGE_OK = 0
KGE_ERR_INVALID_ARGUMENT = -1
KGE_ERR_WORKSPACE = -2
KGE_ERR_CUDA = -3
KGE_ERR_UNSUPPORTED = -4


class ModelDesc(C.Structure):
    _fields_ = [
        ("model", C.c_int32), ("hidden_dim", C.c_int32), ("entity_dim", C.c_int32), ("relation_dim", C.c_int32),
        ("nentity", C.c_int64), ("nrelation", C.c_int64),
        ("gamma", C.c_float), ("embedding_range", C.c_float), ("interht_eps", C.c_float), ("reserved", C.c_int32),
    ]


class LossOpts(C.Structure):
    _fields_ = [
        ("negative_adversarial_sampling", C.c_int32), ("uni_weight", C.c_int32), ("detach_weights", C.c_int32),
        ("reserved", C.c_int32), ("adversarial_temperature", C.c_float), ("regularization", C.c_float),
    ]


class AdamOpts(C.Structure):
    _fields_ = [
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("step", C.c_int32), ("keras_epsilon_mode", C.c_int32),
    ]


class TranSparseDesc(C.Structure):
    _fields_ = [("nentity", C.c_int64), ("nrelation", C.c_int64), ("dim", C.c_int64), ("gamma", C.c_float),
                ("project_tail", C.c_int32)]


class StepBuffers(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("S", "Q", "GH", "GT", "GR", "GM", "stats", "terms")]


_P = C.c_void_p
_I64 = C.c_int64
_SZ = C.c_size_t
_DESC = C.POINTER(ModelDesc)

# name -> (restype, argtypes): exactly the prototypes of include/kge_b200.h
PROTOTYPES = {
    "kge_last_error": (C.c_char_p, []),
    "kge_abi_version": (C.c_int, []),
    "kge_device_info": (C.c_int, [C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_char_p, _SZ]),
    "kge_score_forward": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _P, _I64, _I64, _P, _P]),
    "kge_score_backward": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _P, _I64, _I64, _P, _P, _P, _P, C.c_int, _P, _SZ, _P]),
    "kge_loss_forward_backward": (C.c_int, [_P, _P, _P, _I64, _I64, C.POINTER(LossOpts), _P, _P, _P, _P, _SZ, _P]),
    "kge_loss_static_max_b": (C.c_int64, []),
    "kge_loss_forward_backward_static": (C.c_int, [_P, _P, _P, _I64, _I64, C.POINTER(LossOpts), _P, _P, _P, _P]),
    "kge_l3_sum_cubes": (C.c_int, [_P, _I64, _P, _P, _SZ, _P]),
    "kge_adam_dense": (C.c_int, [_P, _P, _P, _P, _I64, C.POINTER(AdamOpts), C.c_float, C.c_int, _P]),
    "kge_eval_rank": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P, _P, _P, C.c_int, _P, _SZ, _P]),
    "kge_eval_rank_shard": (C.c_int, [_DESC, C.c_int, _P, _I64, _I64, _P, _P, _P, _P, _I64, _I64, _I64, _P, _P, _P, _P, _P,
                                      C.c_int, _P, _SZ, _P]),
    "kge_sample_negatives": (C.c_int, [_P, _I64, _I64, C.c_int, _I64, _P, _I64, C.c_uint64, C.c_uint64, C.c_uint32, _P, _P]),
    "kge_crc32c": (C.c_uint32, [_P, _SZ]),
    "kge_tfrecord_next": (C.c_int, [_P, _SZ, _SZ, C.POINTER(_SZ), C.POINTER(_SZ), C.POINTER(_SZ), C.c_int]),
    "kge_tfrecord_frame": (_SZ, [_P, _SZ, _P]),
    "kge_record_parse_example": (C.c_int, [_P, _SZ, _P, _SZ, _P, _SZ, _P, _SZ, _P, _SZ, C.POINTER(_SZ)]),
    "kge_record_build_example": (_SZ, [_P, _SZ, _P, _SZ, _P, _SZ, _P, _SZ, _P, _SZ]),
    "kge_eval_set_tensor_path": (C.c_int, [C.c_int]),
    "kge_train_set_sweep_variant": (C.c_int, [C.c_int]),
    "kge_train_step": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64,
                                 C.POINTER(LossOpts), C.POINTER(AdamOpts), C.c_int, _P, _P, _P, _SZ, _P]),
    "kge_train_step_host": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64,
                                      C.POINTER(LossOpts), C.POINTER(AdamOpts), C.c_int, _P, _P, _SZ, _P]),
    "kge_stage_batch": (C.c_int, [_P, _P, _P, _P, _P, _P, _I64, _I64, _P]),
    "kge_fetch_floats": (C.c_int, [_P, _P, _I64, _P]),
    "kge_stream_wait_event": (C.c_int, [_P, _P]),
    "kge_train_sweep": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _P, _P, _I64, _I64, C.POINTER(LossOpts), _P, _I64,
                                  C.POINTER(StepBuffers), _P, _SZ, _P]),
    "kge_train_apply": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64,
                                  C.POINTER(LossOpts), C.POINTER(AdamOpts), C.POINTER(StepBuffers), _I64, _I64, _P, _P,
                                  _P, _SZ, _P]),
    "kge_train_apply_peers": (C.c_int, [_DESC, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64,
                                        C.POINTER(LossOpts), C.POINTER(AdamOpts), C.POINTER(StepBuffers), _I64, _I64,
                                        C.c_int, C.POINTER(C.c_void_p), _P, _P, _P, _P, _SZ, _P]),
    "kge_train_exchange": (C.c_int, [_DESC, _P, _P, _I64, _I64, C.POINTER(StepBuffers), C.c_int, C.c_int,
                                     C.POINTER(StepBuffers), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), _P]),
    "kge_train_apply_shard": (C.c_int, [_DESC, C.c_int, _P, _I64, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64,
                                        C.POINTER(LossOpts), C.POINTER(AdamOpts), C.POINTER(StepBuffers), _I64, _I64, _P,
                                        _P, _P, _SZ, _P]),
    "kge_shard_compact_workspace_bytes": (_SZ, [_I64]),
    "kge_shard_compact": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _P, _P, _I64, _P, _P, _SZ, _P]),
    "kge_shard_fetch_rows": (C.c_int, [_P, _P, _I64, _I64, C.POINTER(C.c_void_p), C.c_int, _I64, _P, _P]),
    "kge_workspace_bytes": (_SZ, [C.c_int, _DESC, _I64, _I64]),
    "kge_transparse_nout": (C.c_int64, [C.POINTER(TranSparseDesc), C.c_int, _I64]),
    "kge_transparse_workspace_bytes": (_SZ, [C.POINTER(TranSparseDesc), C.c_int, _I64, _I64, C.c_int]),
    "kge_transparse_forward": (C.c_int, [C.POINTER(TranSparseDesc), C.c_int, _P, _P, _P, _P, _P, _P, _I64, _I64, _P, _P,
                                         C.c_int, _P, _SZ, _P]),
    "kge_transparse_backward": (C.c_int, [C.POINTER(TranSparseDesc), C.c_int, _P, _P, _P, _I64, _I64, _P, _P, _P, _P, _P,
                                          _P, _SZ, _P]),
    "kge_tf_call_terms": (C.c_int, [_P, _I64, _I64, C.c_int, _P, _P]),
    "kge_tf_call_terms_backward": (C.c_int, [_P, _P, _I64, _I64, C.c_int, _P, _P]),
    "kge_check_ids": (C.c_int, [_P, _P, _I64, _I64, _I64, _I64, _P, _P]),
    "kge_probe_l2_read": (C.c_int, [_P, _SZ, C.c_int, _P, C.POINTER(C.c_int64), _P]),
    "kge_profile_enable": (C.c_int, [C.c_int]),
    "kge_profile_num_stages": (C.c_int, []),
    "kge_profile_stage_name": (C.c_char_p, [C.c_int]),
    "kge_profile_read": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_int64), C.c_int]),
}

_lib = None


class KgeError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load the library once.  Raises if it is absent -- there is no CPU / eager fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise KgeError(
            "libkge_b200.so not found at %s: build it with `python -m customknowledgegraphembedding_b200.build` "
            "(needs nvcc); this package has no fallback path" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (restype, argtypes) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    got = lib.kge_abi_version()
    if got != ABI_VERSION:
        raise KgeError("libkge_b200.so ABI version %d, binding expects %d: rebuild" % (got, ABI_VERSION))
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    """Map a kge_status to the exception the reference raises in the same situation:
    invalid model / mode / dims -> ValueError (KGEModel.__init__/forward); else RuntimeError."""
    if rc == KGE_OK:
        return
    msg = load().kge_last_error().decode("utf-8", "replace")
    if rc == KGE_ERR_INVALID_ARGUMENT:
        raise ValueError(msg)
    raise KgeError("%s failed (%d): %s" % (what or "libkge_b200", rc, msg))
