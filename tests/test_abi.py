"""CPU: the C-ABI library loads and exports every symbol include/kge_b200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(REPO, "include", "kge_b200.h")).read()
    return sorted(set(re.findall(r"KGE_API\s+[\w\s\*]+?\b(kge_\w+)\s*\(", text)))


def test_header_declares_entry_points():
    names = _declared()
    for must in ("kge_score_forward", "kge_score_backward", "kge_loss_forward_backward", "kge_adam_dense",
                 "kge_eval_rank", "kge_train_step", "kge_train_step_host", "kge_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from customknowledgegraphembedding_b200 import _lib, build
    if not os.path.exists(_lib.LIB_PATH):
        build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), "libkge_b200.so does not export %s" % name
        assert name in _lib.PROTOTYPES, "ctypes binding lacks %s" % name
    assert set(_lib.PROTOTYPES) == set(_declared())


def test_binding_loads_and_reports_abi():
    from customknowledgegraphembedding_b200 import _lib
    assert _lib.load().kge_abi_version() == _lib.ABI_VERSION


def test_invalid_desc_is_value_error_without_gpu():
    """Argument validation happens before any CUDA call, so it is checkable on the CPU box."""
    from customknowledgegraphembedding_b200 import _lib, ops
    lib = _lib.load()
    desc = ops.make_desc("RotatE", 10, 2, 8, 8, 8, 9.0, 1.0)  # RotatE needs entity_dim = 2H
    rc = lib.kge_score_forward(ctypes.byref(desc), 0, None, None, None, None, None, 1, 1, None, None)
    assert rc == _lib.KGE_ERR_INVALID_ARGUMENT
    with pytest.raises(ValueError):
        _lib.check(rc)
    assert ops.workspace_bytes(_lib.WS_TRAIN_STEP, ops.make_desc("TransE", 100, 3, 8, 8, 8, 12.0, 1.75), 4, 8) > 0


def test_no_cpu_fallback():
    import torch
    import customknowledgegraphembedding_b200 as kb
    m = kb.KGEModel("TransE", 10, 2, 8, 12.0)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError):
            m(torch.zeros(2, 3, dtype=torch.int64))
    # the product never imports the oracle
    import sys
    src = open(os.path.join(REPO, "customknowledgegraphembedding_b200", "model.py")).read()
    assert "oracle" not in src and "oracle" not in open(os.path.join(REPO, "customknowledgegraphembedding_b200", "ops.py")).read()


def test_transparse_host_logic_without_gpu():
    """Sizes, the output shape rule and argument validation of the TranSparse entry points need no device."""
    from customknowledgegraphembedding_b200 import _lib
    lib = _lib.load()
    as_written = _lib.TranSparseDesc(100, 7, 48, 12.0, 0)
    projected = _lib.TranSparseDesc(100, 7, 48, 12.0, 1)
    assert lib.kge_transparse_nout(ctypes.byref(as_written), _lib.MODE_IDS["head-batch"], 9) == 9
    assert lib.kge_transparse_nout(ctypes.byref(as_written), _lib.MODE_IDS["tail-batch"], 9) == 1   # model.py:229
    assert lib.kge_transparse_nout(ctypes.byref(projected), _lib.MODE_IDS["tail-batch"], 9) == 9
    assert lib.kge_transparse_nout(ctypes.byref(as_written), _lib.MODE_IDS["single"], 9) == 0
    fwd = lib.kge_transparse_workspace_bytes(ctypes.byref(as_written), 1, 8, 16, 0)
    bwd = lib.kge_transparse_workspace_bytes(ctypes.byref(as_written), 1, 8, 16, 1)
    assert 0 < fwd < bwd
    assert lib.kge_transparse_workspace_bytes(ctypes.byref(_lib.TranSparseDesc(100, 0, 48, 12.0, 0)), 1, 8, 16, 0) == 0
    rc = lib.kge_transparse_forward(ctypes.byref(as_written), 1, None, None, None, None, None, None, 8, 16, None, None, 0,
                                    None, 0, None)
    assert rc == _lib.KGE_ERR_INVALID_ARGUMENT
    rc = lib.kge_transparse_forward(ctypes.byref(as_written), 7, None, None, None, None, None, None, 8, 16, None, None, 0,
                                    None, 0, None)
    assert rc == _lib.KGE_ERR_INVALID_ARGUMENT and b"mode" in lib.kge_last_error()


def test_struct_layouts_match_the_header(tmp_path):
    """The header compiles as plain C (the boundary is a C ABI) and every struct of the ctypes binding has the size and
    field offsets the C compiler gives it."""
    import shutil
    import subprocess
    from customknowledgegraphembedding_b200 import _lib
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    structs = {"kge_model_desc": _lib.ModelDesc, "kge_loss_opts": _lib.LossOpts, "kge_adam_opts": _lib.AdamOpts,
               "kge_step_buffers": _lib.StepBuffers, "kge_transparse_desc": _lib.TranSparseDesc}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "kge_b200.h"', 'int main(void) {']
    for cname, ct in structs.items():
        lines.append('printf("%s size %%zu\\n", sizeof(%s));' % (cname, cname))
        for fname, _ in ct._fields_:
            lines.append('printf("%s %s %%zu\\n", offsetof(%s, %s));' % (cname, fname, cname, fname))
    lines += ['printf("abi %d\\n", KGE_B200_ABI_VERSION);', 'return 0; }']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(REPO, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split("\n")
    seen = {}
    for ln in out:
        parts = ln.split()
        if len(parts) == 3:
            seen[(parts[0], parts[1])] = int(parts[2])
        elif len(parts) == 2 and parts[0] == "abi":
            assert int(parts[1]) == _lib.ABI_VERSION
    for cname, ct in structs.items():
        assert seen[(cname, "size")] == ctypes.sizeof(ct), cname
        for fname, _ in ct._fields_:
            assert seen[(cname, fname)] == getattr(ct, fname).offset, (cname, fname)


def test_binding_argument_counts_match_the_header():
    """Every prototype of the ctypes binding takes as many arguments as the declaration in include/kge_b200.h."""
    from customknowledgegraphembedding_b200 import _lib
    text = open(os.path.join(REPO, "include", "kge_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    found = {}
    for m in re.finditer(r"KGE_API\s+[\w\s\*]+?\b(kge_\w+)\s*\(([^;]*?)\)\s*;", text, flags=re.S):
        params = m.group(2).strip()
        found[m.group(1)] = 0 if params in ("", "void") else params.count(",") + 1
    assert set(found) == set(_lib.PROTOTYPES)
    for name, (_, argtypes) in _lib.PROTOTYPES.items():
        assert found[name] == len(argtypes), "%s: header has %d parameters, binding %d" % (name, found[name], len(argtypes))


def test_transparse_has_no_cpu_path_either():
    import torch
    from customknowledgegraphembedding_b200.transparse import TranSparseModel
    m = TranSparseModel(10, 2, 8, 12.0)
    assert m.W.shape == (2, 8, 8) and m.mask.shape == (2, 8, 8) and set(m.mask.unique().tolist()) <= {0.0, 1.0}
    assert [n for n, p in m.named_parameters() if p.requires_grad] == ["entity_embedding", "relation_embedding", "W"]
    assert m.nout("head-batch", 7) == 7 and m.nout("tail-batch", 7) == 1 and m.nout("single", 7) == 0
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            m(torch.zeros(2, 3, dtype=torch.int64))
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            m((torch.zeros(2, 3, dtype=torch.int64), torch.zeros(2, 4, dtype=torch.int64)), "head-batch")
    # the product never imports the oracle
    src = open(os.path.join(REPO, "customknowledgegraphembedding_b200", "transparse.py")).read()
    assert "oracle" not in src
