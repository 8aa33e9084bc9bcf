"""kge_loss_forward_backward_static -- K2 with its scratch in a file-scope __device__ array (no allocation, no caller
workspace; include/kge_b200.h) -- against the existing kge_loss_forward_backward, through the C-ABI, on two calls in a
row.  The two entry points launch the same two kernels, so the outputs are required to be EQUAL (which is inside the
1e-5 the loss tests use); the second call runs a different, smaller batch so that it meets the first call's leftovers
in the shared array.

No counterpart in the reference.  This file is
collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(B, N, seed, adv, uni, detach, alpha):
    g = torch.Generator().manual_seed(seed)
    neg = (torch.randn(B, N, generator=g) * 4).cuda()
    pos = (torch.randn(B, generator=g) * 4).cuda()
    w = torch.sqrt(1.0 / torch.randint(8, 201, (B,), generator=g).float()).cuda()
    return neg, pos, w, (adv, uni, detach, alpha)


def _run_both(lib, _lib, neg, pos, w, opts):
    from customknowledgegraphembedding_b200._lib import LossOpts
    B, N = neg.shape
    adv, uni, detach, alpha = opts
    lo = LossOpts(int(adv), int(uni), int(detach), 0, float(alpha), 0.0)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    P = lambda t: C.c_void_p(t.data_ptr())
    out = {}
    for name in ("workspace", "static"):
        # outputs are the caller's (allocated here, before the call); poisoned so that an unwritten element shows
        loss3 = torch.full((3,), float("nan"), device="cuda")
        dneg = torch.full((B, N), float("nan"), device="cuda")
        dpos = torch.full((B,), float("nan"), device="cuda")
        if name == "workspace":
            ws = torch.empty(2 * B + 8, dtype=torch.float32, device="cuda")
            rc = lib.kge_loss_forward_backward(P(neg), P(pos), None if uni else P(w), B, N, C.byref(lo), P(loss3), P(dneg),
                                               P(dpos), P(ws), ws.numel() * 4, stream)
        else:
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
            rc = lib.kge_loss_forward_backward_static(P(neg), P(pos), None if uni else P(w), B, N, C.byref(lo), P(loss3),
                                                      P(dneg), P(dpos), stream)
            torch.cuda.synchronize()
            assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before  # torch's allocator untouched
        _lib.check(rc, name)
        out[name] = (loss3, dneg, dpos)
    torch.cuda.synchronize()
    return out


def test_static_scratch_loss_equals_workspace_loss_twice():
    from customknowledgegraphembedding_b200 import _lib
    lib = _lib.load()
    assert lib.kge_loss_static_max_b() == 65536
    # first call: a large batch fills the shared array; second call: a smaller, different batch right behind it
    cases = [_case(1024, 256, 1, True, False, True, 1.0), _case(37, 129, 2, True, False, False, 0.5),
             _case(512, 64, 3, False, True, True, 1.0)]
    for neg, pos, w, opts in cases:
        out = _run_both(lib, _lib, neg, pos, w, opts)
        for a, b, what in zip(out["workspace"], out["static"], ("loss3", "dneg", "dpos")):
            assert torch.isfinite(b).all(), what
            assert torch.equal(a, b), "%s differs: max |diff| %.3e" % (what, float((a - b).abs().max()))
            # (equal is inside the tolerance the loss tests use: 1e-5 relative)
            torch.testing.assert_close(b, a, rtol=1e-5, atol=0.0)
    # and against the value the first case must have whatever ran in between (the array keeps old contents)
    neg, pos, w, opts = cases[0]
    again = _run_both(lib, _lib, neg, pos, w, opts)
    first = _run_both(lib, _lib, neg, pos, w, opts)
    assert torch.equal(again["static"][0], first["static"][0]) and torch.equal(again["static"][1], first["static"][1])


def test_static_scratch_limit_is_enforced():
    from customknowledgegraphembedding_b200 import _lib
    from customknowledgegraphembedding_b200._lib import LossOpts
    lib = _lib.load()
    B = lib.kge_loss_static_max_b() + 1
    neg = torch.zeros(B, 1, device="cuda")
    pos = torch.zeros(B, device="cuda")
    loss3 = torch.zeros(3, device="cuda")
    lo = LossOpts(0, 1, 1, 0, 1.0, 0.0)
    P = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.kge_loss_forward_backward_static(P(neg), P(pos), None, B, 1, C.byref(lo), P(loss3), None, None,
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.kge_last_error().decode()
    assert "KGE_LOSS_STATIC_MAX_B" in msg and str(B) in msg, msg
    with pytest.raises(ValueError):
        _lib.check(rc, "kge_loss_forward_backward_static")
    # the largest accepted size works
    B = lib.kge_loss_static_max_b()
    rc = lib.kge_loss_forward_backward_static(P(neg), P(pos), None, B, 1, C.byref(lo), P(loss3), None, None,
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, "kge_loss_forward_backward_static")
    torch.cuda.synchronize()
    # loss = (-logsigmoid(0) - logsigmoid(-0)) / 2 = log 2 (65536 fp32 terms: the loss tests' 1e-5 relative), and equal
    # to the existing entry point's
    assert abs(loss3[2].item() - 0.6931472) < 1e-5 * 0.6931472
    want = torch.zeros(3, device="cuda")
    ws = torch.empty(2 * B + 8, dtype=torch.float32, device="cuda")
    _lib.check(lib.kge_loss_forward_backward(P(neg), P(pos), None, B, 1, C.byref(lo), P(want), None, None, P(ws), ws.numel() * 4,
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)), "kge_loss_forward_backward")
    torch.cuda.synchronize()
    assert torch.equal(want, loss3)
