"""GPU: the multi-rank split of the step (kge_train_sweep + gather + kge_train_apply) equals the single call,
emulated on one GPU by sweeping the two halves of a batch separately (never two processes on one GPU)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import kge_oracle as ko

pytestmark = pytest.mark.gpu
GAMMA = {"TransE": 12.0, "DistMult": 12.0, "ComplEx": 12.0, "RotatE": 9.0, "pRotatE": 9.0, "InterHT": 24.0}


def _model(name, E, R, H, seed):
    import customknowledgegraphembedding_b200 as kb
    om = ko.make_model(name, E, R, H, GAMMA[name], seed=seed)
    de, dr, tr = ko.model_flags(name)
    gm = kb.KGEModel(name, E, R, H, GAMMA[name], de, dr, tr)
    gm.load_state_dict(om.state_dict())
    return gm.cuda()


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
@pytest.mark.parametrize("mode", ["head-batch", "tail-batch"])
def test_two_sweeps_plus_apply_equal_one_step(name, mode):
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H, B, N, lr = 200, 5, 64, 16, 24, 1e-3
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=0.5, cuda=True)
    pos, neg, w, _ = ko.synthetic_batch(E, R, B, N, seed=1)
    pos, neg, w = pos.cuda(), neg.cuda(), w.cuda()

    ref = _model(name, E, R, H, seed=2)
    ref_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, ref.parameters()), lr=lr)
    want = kb.KGEModel.train_step(ref, ref_opt, iter([(pos, neg, w, mode)]), args)

    gm = _model(name, E, R, H, seed=2)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=lr)
    step = kd.DataParallelStep(gm, opt)           # world 1: we drive its two halves by hand
    half = B // 2
    loss_opts = kd.LossOpts(1, 0, 1, 0, 0.5, 0.0)
    wsum = w.sum().reshape(1)
    gathered = {}
    for r in range(2):
        sl = slice(r * half, (r + 1) * half)
        step._shape = None
        step.world = 1
        step._ensure(half, N, pos.device)
        step._sweep(mode, pos[sl].contiguous(), neg[sl].contiguous(), w[sl].contiguous(), wsum, B, loss_opts)
        for k, v in step._local.items():
            gathered.setdefault(k, []).append(v.clone())
    step._gathered = {k: torch.cat(v) for k, v in gathered.items()}
    for p in (gm.entity_embedding, gm.relation_embedding) + ((gm.modulus,) if name == "pRotatE" else ()):
        st = gm._adam_state(opt, p)
        st["step"] += 1
    adam = kd.AdamOpts(lr, 0.9, 0.999, 1e-8, 1, 0)
    out = torch.zeros(4, device="cuda")
    # apply in two row ranges would update the relation table twice, so: one apply over all rows
    step.lo, step.hi = 0, E
    step._apply(mode, pos, neg, loss_opts, adam, out)
    got = out.tolist()
    np.testing.assert_allclose(got[2], want["loss"], rtol=2e-6)
    for pn in ("entity_embedding", "relation_embedding"):
        a, b = getattr(gm, pn).detach(), getattr(ref, pn).detach()
        # identical arithmetic except the weight sum (torch reduction vs in-kernel reduction order)
        assert float((a - b).abs().max()) <= 2e-3 * lr, pn


def test_apply_touches_only_owned_rows():
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H, B, N = 100, 3, 32, 8, 12
    gm = _model("RotatE", E, R, H, seed=5)
    before = gm.entity_embedding.detach().clone()
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=1e-2)
    step = kd.DataParallelStep(gm, opt)
    step.lo, step.hi = 30, 70
    batch = ko.synthetic_batch(E, R, B, N, seed=3, mode="tail-batch")
    step.train_step(iter([batch]), ko.Args(negative_adversarial_sampling=True))
    after = gm.entity_embedding.detach()
    assert torch.equal(after[:30], before[:30]) and torch.equal(after[70:], before[70:])
    assert not torch.equal(after[30:70], before[30:70])
    m = opt.state[gm.entity_embedding]["exp_avg"]
    assert float(m[:30].abs().max()) == 0.0 and float(m[70:].abs().max()) == 0.0


def test_world1_step_equals_train_step():
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H, B, N = 150, 4, 32, 8, 10
    a, b = _model("TransE", E, R, H, seed=7), _model("TransE", E, R, H, seed=7)
    oa = torch.optim.Adam(filter(lambda p: p.requires_grad, a.parameters()), lr=1e-3)
    ob = torch.optim.Adam(filter(lambda p: p.requires_grad, b.parameters()), lr=1e-3)
    args = ko.Args(negative_adversarial_sampling=False, uni_weight=True, cuda=True)
    step = kd.DataParallelStep(b, ob)
    for s in range(3):
        batch = ko.synthetic_batch(E, R, B, N, seed=s)
        la = kb.KGEModel.train_step(a, oa, iter([batch]), args)
        lb = step.train_step(iter([batch]), args)
        assert la["loss"] == lb["loss"]
    assert torch.equal(a.entity_embedding.detach(), b.entity_embedding.detach())
    assert torch.equal(a.relation_embedding.detach(), b.relation_embedding.detach())


def test_sharded_eval_counts_add_up():
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H = 333, 4, 32
    gm = _model("RotatE", E, R, H, seed=1)
    rng = np.random.RandomState(0)
    q = torch.from_numpy(np.stack([rng.randint(0, E, 50), rng.randint(0, R, 50), rng.randint(0, E, 50)], 1))
    _, ng, ne = gm.rank_queries(q, "tail-batch")
    tot_g, tot_e = torch.zeros_like(ng), torch.zeros_like(ne)
    for r in range(3):
        lo, hi = kd.shard_range(E, r, 3)
        _, g, e = gm.rank_queries(q, "tail-batch", entity_lo=lo, entity_hi=hi)
        tot_g += g
        tot_e += e
    assert torch.equal(tot_g, ng) and torch.equal(tot_e, ne)
