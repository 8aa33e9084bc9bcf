"""GPU parity: the sm_100a kernels (through the C-ABI) against the CPU oracle on the same seeded inputs.

Tolerances: BASELINE.json north_star -- scores and loss within 1e-5 relative in fp32, per-step embedding
updates within 1e-5.  Scores are gamma - sum(...) so the absolute term is scaled by the magnitude of the
quantity being summed (|gamma| + |score|)."""
import os

import numpy as np
import pytest
import torch

from oracle import kge_oracle as ko

pytestmark = pytest.mark.gpu

GAMMA = {"TransE": 12.0, "DistMult": 12.0, "ComplEx": 12.0, "RotatE": 9.0, "pRotatE": 9.0, "InterHT": 24.0}
RTOL = 1e-5


def _pair(name, E, R, H, seed=0):
    """(oracle model on CPU, our KGEModel on cuda:0) with identical weights."""
    import customknowledgegraphembedding_b200 as kb
    om = ko.make_model(name, E, R, H, GAMMA[name], seed=seed)
    de, dr, tr = ko.model_flags(name)
    gm = kb.KGEModel(name, E, R, H, GAMMA[name], de, dr, tr)
    gm.load_state_dict(om.state_dict())
    return om, gm.cuda()


def _close(got, want, scale, rtol=RTOL, what="", atol=0.0):
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    err = np.abs(got - want)
    tol = rtol * (np.abs(want) + scale) + atol
    bad = err > tol
    assert not bad.any(), "%s: %d/%d outside 1e-5 (max err %.3e, tol there %.3e)" % (
        what, bad.sum(), bad.size, err.max(), tol.reshape(-1)[err.argmax()])


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
@pytest.mark.parametrize("H", [32, 500, 30])  # 30: not a multiple of 4 -> scalar path
def test_forward_scores(name, H):
    E, R, B, N = 211, 7, 9, 37
    om, gm = _pair(name, E, R, H)
    pos, neg, _, _ = ko.synthetic_batch(E, R, B, N, seed=3)
    for mode in ko.MODES:
        sample = pos if mode == "single" else (pos, neg)
        with torch.no_grad():
            want = om(sample, mode).numpy()
            got = gm(pos.cuda() if mode == "single" else (pos.cuda(), neg.cuda()), mode).cpu().numpy()
        assert got.shape == want.shape
        # the summed magnitude is gamma - score for distance models
        _close(got, want, scale=GAMMA[name] + np.abs(want).max(), what="%s %s H=%d" % (name, mode, H))


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
@pytest.mark.parametrize("deterministic", [False, True])
@pytest.mark.parametrize("H", [32, 30])
def test_backward_gradients(name, deterministic, H):
    E, R, B, N = 97, 5, 6, 19
    om, gm = _pair(name, E, R, H, seed=5)
    gm.deterministic = deterministic
    pos, neg, _, _ = ko.synthetic_batch(E, R, B, N, seed=9)
    neg[0, :3] = neg[0, 0]          # repeated negatives inside one positive
    neg[1, 0] = pos[1, 2]           # a negative equal to the positive's own tail
    g = torch.Generator().manual_seed(1)
    for mode in ko.MODES:
        om.zero_grad()
        gm.zero_grad()
        sample = pos if mode == "single" else (pos, neg)
        s_ref = om(sample, mode)
        up = torch.randn(s_ref.shape, generator=g)
        s_ref.backward(up)
        s = gm(pos.cuda() if mode == "single" else (pos.cuda(), neg.cuda()), mode)
        s.backward(up.cuda())
        for pname in ("entity_embedding", "relation_embedding") + (("modulus",) if name == "pRotatE" else ()):
            want = getattr(om, pname).grad.numpy()
            got = getattr(gm, pname).grad.cpu().numpy()
            scale = np.abs(want).max()
            _close(got, want, scale=scale, rtol=2e-5, what="%s %s grad %s det=%s" % (name, mode, pname, deterministic))


def test_backward_deterministic_is_bit_reproducible():
    E, R, B, N, H = 300, 5, 32, 64, 64
    _, gm = _pair("RotatE", E, R, H)
    gm.deterministic = True
    pos, neg, _, _ = ko.synthetic_batch(E, R, B, N, seed=2)
    pos, neg = pos.cuda(), neg.cuda()
    up = torch.randn(B, N, device="cuda")
    grads = []
    for _ in range(3):
        gm.zero_grad()
        gm((pos, neg), "tail-batch").backward(up)
        grads.append(gm.entity_embedding.grad.clone())
    assert torch.equal(grads[0], grads[1]) and torch.equal(grads[0], grads[2])


@pytest.mark.parametrize("adv,uni,detach,alpha", [(True, False, True, 1.0), (True, False, True, 0.5),
                                                  (True, True, True, 1.0), (False, False, True, 1.0),
                                                  (False, True, True, 1.0), (True, False, False, 1.0)])
def test_loss_kernel(adv, uni, detach, alpha):
    from customknowledgegraphembedding_b200 import ops
    B, N = 33, 257
    g = torch.Generator().manual_seed(4)
    neg = (torch.randn(B, N, generator=g) * 4).requires_grad_(True)
    pos = (torch.randn(B, 1, generator=g) * 4).requires_grad_(True)
    w = torch.sqrt(1.0 / torch.randint(8, 201, (B,), generator=g).float())
    pl, nl, loss = ko.kge_loss(neg, pos, w, negative_adversarial_sampling=adv, adversarial_temperature=alpha,
                               uni_weight=uni, detach_weights=detach)
    loss.backward()
    loss3, dneg, dpos = ops.loss_forward_backward(neg.detach().cuda(), pos.detach().cuda(), w.cuda(),
                                                  negative_adversarial_sampling=adv, adversarial_temperature=alpha,
                                                  uni_weight=uni, detach_weights=detach)
    _close(loss3.cpu().numpy(), [pl.item(), nl.item(), loss.item()], scale=0.0, what="loss3")
    _close(dneg.cpu().numpy(), neg.grad.numpy(), scale=np.abs(neg.grad.numpy()).max(), what="dneg")
    _close(dpos.cpu().numpy(), pos.grad.numpy().reshape(-1), scale=np.abs(pos.grad.numpy()).max(), what="dpos")


def test_loss_golden_reference_source():
    """K2 against the values the reference's own loss() produced (tests/golden/ref_tf_loss.npz)."""
    from customknowledgegraphembedding_b200 import ops
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_tf_loss.npz"))
    neg = torch.from_numpy(g["negative_score"]).cuda()
    pos = torch.from_numpy(g["positive_score"]).cuda()
    w = torch.from_numpy(g["subsampling_weight"]).cuda()
    reg_e = ops.l3_sum_cubes(torch.from_numpy(g["entity_embedding"]).cuda().reshape(-1)).item()
    reg_r = ops.l3_sum_cubes(torch.from_numpy(g["relation_embedding"]).cuda().reshape(-1)).item()
    for adv, uni, reg, temp, want in g["cases"]:
        loss3, _, _ = ops.loss_forward_backward(neg, pos, w, negative_adversarial_sampling=bool(adv),
                                                adversarial_temperature=float(temp), uni_weight=bool(uni),
                                                want_grad=False)
        got = loss3[2].item() + float(reg) * (reg_e + reg_r)
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5)


def test_adam_kernel_matches_torch_adam():
    from customknowledgegraphembedding_b200 import ops
    n = 4099  # not a multiple of 4: exercises the tail
    g = torch.Generator().manual_seed(0)
    p0 = torch.randn(n, generator=g)
    p_ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=1e-2)
    p = p0.clone().cuda()
    m = torch.zeros_like(p)
    v = torch.zeros_like(p)
    for step in range(1, 6):
        grad = torch.randn(n, generator=g) * (0.0 if step == 3 else 1.0)  # an all-zero gradient step too
        p_ref.grad = grad.clone()
        opt.step()
        gg = grad.cuda()
        ops.adam_dense(p, gg, m, v, lr=1e-2, step=step, zero_grad=True)
        assert float(gg.abs().max()) == 0.0
        delta_ref = (p_ref.detach() - p0).numpy()
        # p ~ 1: one fp32 ulp of p (1.2e-7) bounds how exactly any update can be observed
        _close((p.cpu() - p0).numpy(), delta_ref, scale=1e-2, rtol=2e-5, what="adam step %d" % step)


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
@pytest.mark.parametrize("host_batch", [True, False])
@pytest.mark.parametrize("deterministic", [0, 2, 3])  # step variant: fused (default), v1 atomic, v1 fixed-order
def test_train_step_matches_oracle(name, host_batch, deterministic):
    """3 consecutive steps (head/tail alternating, -adv, subsampling weights): loss within 1e-5 and the
    per-step update of both tables within 1e-5 of the oracle's torch.optim.Adam step."""
    E, R, H, B, N, lr = 151, 6, 32, 16, 24, 1e-3
    om, gm = _pair(name, E, R, H, seed=7)
    gm.deterministic = deterministic
    reg = 1e-4 if name in ("DistMult", "ComplEx") else 0.0
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=0.5, regularization=reg, cuda=True)
    o_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, om.parameters()), lr=lr)
    g_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=lr)
    for step in range(3):
        pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=20 + step)
        before = om.entity_embedding.detach().clone(), om.relation_embedding.detach().clone()
        want = ko.train_step(om, o_opt, iter([(pos, neg, w, mode)]), args)
        batch = (pos, neg, w, mode) if host_batch else (pos.cuda(), neg.cuda(), w.cuda(), mode)
        got = type(gm).train_step(gm, g_opt, iter([batch]), args)
        assert set(got) == set(want)
        for k in want:
            np.testing.assert_allclose(got[k], want[k], rtol=1e-5, atol=1e-6, err_msg="%s step %d %s" % (name, step, k))
        for pname, b4 in zip(("entity_embedding", "relation_embedding"), before):
            upd_ref = (getattr(om, pname).detach() - b4).numpy()
            upd = (getattr(gm, pname).detach().cpu() - b4).numpy()
            # Adam's update is ~lr * sign(g) on first touch: compare updates against lr.  An update is only
            # observable to one fp32 ulp of the parameter it is applied to (p_after - p_before), hence atol.
            ulp = 2.0 * float(b4.abs().max()) * 2.0 ** -23  # both sides round p_after: up to 1 ulp apart + the real error
            _close(upd, upd_ref, scale=lr, rtol=1e-4 if step else 1e-5, atol=ulp,
                   what="%s update %s step %d" % (name, pname, step))
        if name == "pRotatE":
            np.testing.assert_allclose(gm.modulus.item(), om.modulus.item(), rtol=1e-5)
    # the torch optimizer object stays usable / inspectable (same state layout as torch.optim.Adam)
    st = g_opt.state[gm.entity_embedding]
    assert int(st["step"]) == 3 and st["exp_avg"].shape == gm.entity_embedding.shape


@pytest.mark.parametrize("name", ["RotatE", "TransE", "InterHT", "pRotatE"])
@pytest.mark.parametrize("H,N,mode", [(1000, 256, "tail-batch"), (500, 37, "head-batch"), (1500, 19, "head-batch"),
                                      (2500, 9, "tail-batch"), (8, 5, "head-batch")])
def test_fused_step_shapes(name, H, N, mode):
    """The second-generation step at the BASELINE hidden sizes, every chunking (H <= 1024 / 2048 / 4096),
    N not a multiple of the row group, uniform and adversarial weights."""
    E, R, B, lr = 300, 5, 12, 1e-3
    om, gm = _pair(name, E, R, H, seed=11)
    for adv, uni in ((True, False), (False, True)):
        args = ko.Args(negative_adversarial_sampling=adv, adversarial_temperature=1.0, uni_weight=uni, cuda=True)
        o_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, om.parameters()), lr=lr)
        g_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=lr)
        pos, neg, w, _ = ko.synthetic_batch(E, R, B, N, seed=5)
        neg[0, :2] = pos[0, 0]  # candidates that are also the positive's own head
        before = om.entity_embedding.detach().clone()
        want = ko.train_step(om, o_opt, iter([(pos, neg, w, mode)]), args)
        got = type(gm).train_step(gm, g_opt, iter([(pos, neg, w, mode)]), args)
        for k in want:
            np.testing.assert_allclose(got[k], want[k], rtol=1e-5, atol=1e-6, err_msg="%s %s" % (name, k))
        # gradients (Adam's first moment after one step is 0.1 * g) within 1e-5 of the largest entry
        for pn in ("entity_embedding", "relation_embedding"):
            g_ref = o_opt.state[getattr(om, pn)]["exp_avg"].numpy() * 10.0
            g_got = g_opt.state[getattr(gm, pn)]["exp_avg"].cpu().numpy() * 10.0
            # dL/ds carries the softmax weight exp(alpha * s): a score that is right to 1e-6 * |gamma - s| moves
            # the weight by alpha times that, so gradients agree to a few 1e-5 (of the largest entry) at H ~ 1e3
            _close(g_got, g_ref, scale=np.abs(g_ref).max(), rtol=5e-5,
                   what="%s fused grad %s H=%d adv=%s" % (name, pn, H, adv))
        # updates: Adam divides by |g| + 1e-8, so where |g| is within ~1e3 roundings of zero NO two fp32
        # implementations agree to 1e-5; compare the update where the division is well conditioned
        g_ref = o_opt.state[om.entity_embedding]["exp_avg"].numpy() * 10.0
        ok = np.abs(g_ref) > 1e-3 * np.abs(g_ref).max()
        upd_ref = (om.entity_embedding.detach() - before).numpy()
        upd = (gm.entity_embedding.detach().cpu() - before).numpy()
        _close(upd[ok], upd_ref[ok], scale=lr, rtol=5e-5, atol=2.0 * float(before.abs().max()) * 2.0 ** -23,
               what="%s fused update H=%d adv=%s" % (name, H, adv))
        assert np.abs(upd).max() <= lr * (1 + 1e-3)  # |Adam step| <= lr everywhere
        gm.load_state_dict(om.state_dict())  # re-sync for the next configuration


@pytest.mark.parametrize("name", ["RotatE", "ComplEx", "DistMult", "TransE", "pRotatE"])
@pytest.mark.parametrize("H,N,mode", [(1000, 700, "head-batch"), (256, 1, "tail-batch"), (64, 300, "head-batch"),
                                      (1000, 3, "tail-batch"), (512, 129, "tail-batch")])
def test_split_sweep_part_shapes(name, H, N, mode):
    """The warp-per-row sweep at its part boundaries: N > 256 (more than 32 rows per part), N = 1 and N = 3 (fewer
    parts than warps), N = 129 (first size with 8 parts), every row-slot size (H <= 256 / 512 / 1024)."""
    test_fused_step_shapes(name, H, N, mode)


def test_fused_step_is_bit_reproducible():
    E, R, H, B, N = 400, 6, 128, 64, 96
    results = []
    for _ in range(2):
        _, gm = _pair("RotatE", E, R, H, seed=4)
        opt = torch.optim.Adam(gm.parameters(), lr=1e-3)  # gamma / embedding_range ride along, as in upstream run.py
        assert gm._fused_adam_plan(opt) is not None
        args = ko.Args(negative_adversarial_sampling=True, cuda=True)
        for step in range(3):
            batch = ko.synthetic_batch(E, R, B, N, seed=step)
            type(gm).train_step(gm, opt, iter([batch]), args)
        results.append(gm.entity_embedding.detach().clone())
    assert torch.equal(results[0], results[1])


def test_train_step_generic_optimizer_path():
    """A non-Adam optimizer goes through K1/K2/K3 + optimizer.step()."""
    E, R, H, B, N = 80, 4, 16, 8, 12
    om, gm = _pair("TransE", E, R, H, seed=1)
    args = ko.Args(negative_adversarial_sampling=False)
    o_opt = torch.optim.SGD(om.parameters(), lr=0.1)
    g_opt = torch.optim.SGD(gm.parameters(), lr=0.1)
    pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=3)
    want = ko.train_step(om, o_opt, iter([(pos, neg, w, mode)]), args)
    got = type(gm).train_step(gm, g_opt, iter([(pos, neg, w, mode)]), args)
    np.testing.assert_allclose(got["loss"], want["loss"], rtol=1e-5)
    np.testing.assert_allclose(gm.entity_embedding.detach().cpu().numpy(), om.entity_embedding.detach().numpy(),
                               rtol=1e-4, atol=1e-6)


def test_interht_golden_reference_source():
    """K1 against scores produced by the reference's own InterHT source (tests/golden/ref_tf_interht.npz)."""
    import customknowledgegraphembedding_b200 as kb
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_tf_interht.npz"))
    m = kb.KGEModel("InterHT", int(g["E"]), int(g["R"]), int(g["H"]), float(g["gamma"]), True, False, True)
    with torch.no_grad():
        m.entity_embedding.copy_(torch.from_numpy(g["entity_embedding"]))
        m.relation_embedding.copy_(torch.from_numpy(g["relation_embedding"]))
    m = m.cuda()
    pos = torch.from_numpy(g["positive_sample"]).cuda()
    neg = torch.from_numpy(g["negative_sample"]).cuda()
    with torch.no_grad():
        for mode, key in (("single", "score_single"), ("head-batch", "score_head_batch"), ("tail-batch", "score_tail_batch")):
            got = m(pos if mode == "single" else (pos, neg), mode).cpu().numpy()
            _close(got, g[key], scale=float(g["gamma"]), what="InterHT golden " + mode)


def test_errors_like_reference():
    import customknowledgegraphembedding_b200 as kb
    with pytest.raises(ValueError):
        kb.KGEModel("RotatE", 4, 2, 4, 1.0)
    with pytest.raises(ValueError):
        kb.KGEModel("Nope", 4, 2, 4, 1.0)
    m = kb.KGEModel("TransE", 4, 2, 4, 1.0).cuda()
    with pytest.raises(ValueError):
        m(torch.zeros(1, 3, dtype=torch.int64).cuda(), "bad-mode")


@pytest.mark.parametrize("name", ["TransE", "DistMult", "ComplEx", "RotatE", "pRotatE"])
@pytest.mark.parametrize("variant", [0, 1])
def test_sweep_variants(name, variant):
    """Both candidate sweeps of the fused step against the oracle: 1 = warp per row (split sweep, the default where
    it applies), 0 = row groups per CTA (what InterHT / H > 1024 use).  Same tolerances as above."""
    from customknowledgegraphembedding_b200 import _lib
    lib = _lib.load()
    lib.kge_train_set_sweep_variant(variant)
    try:
        test_train_step_matches_oracle(name, False, 0)
        test_fused_step_shapes(name, 1000, 256, "tail-batch")
        test_fused_step_shapes(name, 500, 37, "head-batch")
        test_fused_step_shapes(name, 12, 70, "head-batch")   # several items per positive with a ragged last item
    finally:
        lib.kge_train_set_sweep_variant(1)


def test_sweep_variants_agree():
    """The two sweeps differ in summation order only: same loss to 2e-6, same gradient (Adam's first moment after
    one step is 0.1 * g) to 5e-5 of its largest entry."""
    from customknowledgegraphembedding_b200 import _lib
    lib = _lib.load()
    E, R, H, B, N = 500, 6, 1000, 40, 256
    losses, grads = [], []
    for variant in (0, 1):
        lib.kge_train_set_sweep_variant(variant)
        try:
            _, gm = _pair("RotatE", E, R, H, seed=4)
            opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=1e-3)
            args = ko.Args(negative_adversarial_sampling=True, cuda=True)
            losses.append(type(gm).train_step(gm, opt, iter([ko.synthetic_batch(E, R, B, N, seed=0)]), args)["loss"])
        finally:
            lib.kge_train_set_sweep_variant(1)
        grads.append(opt.state[gm.entity_embedding]["exp_avg"].clone())
    np.testing.assert_allclose(losses[0], losses[1], rtol=2e-6)
    assert float((grads[0] - grads[1]).abs().max()) <= 5e-5 * float(grads[0].abs().max())


def test_train_step_input_lookahead():
    """The input lookahead (batch k+1 taken from the iterator and copied H2D while step k runs) changes neither the
    order in which batches are consumed nor a single bit of the result; StopIteration surfaces on the call that
    upstream's `next(train_iterator)` would raise it on; a second iterator used in between is served in order too."""
    E, R, H, B, N = 300, 5, 64, 32, 40
    args = ko.Args(negative_adversarial_sampling=True, cuda=True)
    batches = [ko.synthetic_batch(E, R, B if s != 3 else B // 2, N, seed=50 + s) for s in range(6)]  # one ragged batch
    batches = [(p.pin_memory(), n.pin_memory(), w.pin_memory(), m) for p, n, w, m in batches]
    runs = []
    for lookahead in (False, True):
        _, gm = _pair("RotatE", E, R, H, seed=9)
        gm.input_lookahead = lookahead
        opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=1e-3)
        it = iter(batches)
        logs = [type(gm).train_step(gm, opt, it, args) for _ in range(len(batches))]
        with pytest.raises(StopIteration):
            type(gm).train_step(gm, opt, it, args)
        runs.append((logs, gm.entity_embedding.detach().clone(), gm.relation_embedding.detach().clone()))
    assert runs[0][0] == runs[1][0]
    assert torch.equal(runs[0][1], runs[1][1]) and torch.equal(runs[0][2], runs[1][2])
    # two iterators interleaved: each is consumed in its own order
    _, gm = _pair("RotatE", E, R, H, seed=9)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=1e-3)
    it_a, it_b = iter(batches[:3]), iter(batches[3:])
    order = [it_a, it_b, it_a, it_b, it_a, it_b]
    _, ref = _pair("RotatE", E, R, H, seed=9)
    ref.input_lookahead = False
    ropt = torch.optim.Adam(filter(lambda p: p.requires_grad, ref.parameters()), lr=1e-3)
    for k, it in enumerate(order):
        got = type(gm).train_step(gm, opt, it, args)
        want = type(ref).train_step(ref, ropt, iter([batches[k // 2 if it is it_a else 3 + k // 2]]), args)
        assert got == want
    assert torch.equal(gm.entity_embedding, ref.entity_embedding)


def test_check_ids_switch_raises_like_index_select():
    """ids are range-checked on request -- on the device for device batches (kge_check_ids), on the host for
    host batches; the message names the table size, the count and an offending id."""
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import ops
    E, R = 50, 4
    pos, neg, w, mode = ko.synthetic_batch(E, R, 8, 16, seed=4)
    ops.check_ids(pos.cuda(), neg.cuda(), E, R)          # in range: returns
    ops.check_ids(pos.cuda(), None, E, R)
    bad_neg = neg.clone()
    bad_neg[3, 5], bad_neg[7, 0] = E, -1
    with pytest.raises(IndexError, match=r"entity id out of range \[0, 50\): 2 ids"):
        ops.check_ids(pos.cuda(), bad_neg.cuda(), E, R)
    bad_pos = pos.clone()
    bad_pos[2, 1] = R + 3
    with pytest.raises(IndexError, match=r"relation id out of range \[0, 4\): 1 ids of the batch, e.g. 7"):
        ops.check_ids(bad_pos.cuda(), neg.cuda(), E, R)
    m = kb.KGEModel("TransE", E, R, 8, 12.0).cuda()
    m.check_ids = True
    m((pos.cuda(), neg.cuda()), "head-batch")
    with pytest.raises(IndexError):
        m((pos.cuda(), bad_neg.cuda()), "head-batch")
    with pytest.raises(IndexError):
        m((bad_pos, neg), "tail-batch")              # host batch: checked on the host
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=1.0, cuda=True)
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    with pytest.raises(IndexError):
        kb.KGEModel.train_step(m, opt, iter([(pos.cuda(), bad_neg.cuda(), w.cuda(), mode)]), args)
