"""GPU parity of K5-T, the tcgen05 (3xTF32, TMA-fed, TMEM-accumulated) path that kge_eval_rank takes for the
GEMM-shaped models DistMult and ComplEx (csrc/kge_eval_mma.cu), against

  * the fp32 CUDA-core tile path of the same entry point (kge_eval_set_tensor_path(0)), and
  * the oracle's upstream-style test_step (tests/test_gpu_eval.py covers that for every model at small sizes;
    with the default path DistMult / ComplEx already run on the tensor cores there).

Ranks are integers; the only legitimate difference is a candidate whose score is within rounding distance of
the true score (different summation order / operand split), bounded per query exactly as in test_gpu_eval.py.
Tolerance on scores: 2e-6 * sum_k |q_k e_k| + 1e-7 (3xTF32 keeps ~22 mantissa bits per product)."""
import numpy as np
import pytest
import torch

from oracle import kge_oracle as ko

pytestmark = pytest.mark.gpu

NAMES = ("DistMult", "ComplEx")


def _model(name, E, R, H, seed=0):
    import customknowledgegraphembedding_b200 as kb
    om = ko.make_model(name, E, R, H, 12.0, seed=seed)
    de, dr, tr = ko.model_flags(name)
    gm = kb.KGEModel(name, E, R, H, 12.0, de, dr, tr)
    gm.load_state_dict(om.state_dict())
    return om, gm.cuda()


def _queries(E, R, Q, seed):
    rng = np.random.RandomState(seed)
    return torch.from_numpy(np.stack([rng.randint(0, E, Q), rng.randint(0, R, Q), rng.randint(0, E, Q)], 1)).cuda()


def _filters(E, Q, seed, max_len=40):
    """Unsorted CSR lists with duplicates; some lists empty, some long."""
    rng = np.random.RandomState(seed)
    lens = rng.randint(0, max_len, Q)
    lens[::7] = 0
    lens[3::11] = 4 * max_len
    indptr = np.zeros(Q + 1, dtype=np.int64)
    np.cumsum(lens, out=indptr[1:])
    idx = rng.randint(0, E, int(indptr[-1])).astype(np.int64)
    return torch.from_numpy(indptr).cuda(), torch.from_numpy(idx).cuda()


class _Path:
    def __init__(self, on):
        self.on = on

    def __enter__(self):
        from customknowledgegraphembedding_b200 import _lib
        _lib.load().kge_eval_set_tensor_path(int(self.on))

    def __exit__(self, *a):
        from customknowledgegraphembedding_b200 import _lib
        _lib.load().kge_eval_set_tensor_path(1)


def _all_scores(gm, queries, mode, E):
    with torch.no_grad():
        cand = torch.arange(E, device="cuda").repeat(queries.shape[0], 1)
        return gm((queries, cand), mode)  # K1, fp32 CUDA cores


def _compare(gm, queries, mode, E, indptr=None, idx=None, lo=0, hi=None):
    hi = E if hi is None else hi
    with _Path(1):
        ts1, ng1, ne1 = gm.rank_queries(queries, mode, indptr, idx, entity_lo=lo, entity_hi=hi)
    with _Path(0):
        ts0, ng0, ne0 = gm.rank_queries(queries, mode, indptr, idx, entity_lo=lo, entity_hi=hi)
    torch.cuda.synchronize()
    s = _all_scores(gm, queries, mode, E)
    col = 0 if mode == "head-batch" else 2
    st = s.gather(1, queries[:, col:col + 1])
    # scale of the rounding noise: |terms| <= |q|.|e| ; use a fixed fraction of max |score| + 1
    tol = 2e-5 * (1.0 + s.abs().max().item())
    assert torch.allclose(ts1, st.view(-1), rtol=0, atol=tol), (ts1 - st.view(-1)).abs().max()
    assert torch.allclose(ts1, ts0, rtol=0, atol=tol)
    near = ((s - st).abs() <= tol).sum(1) - 1
    diff = (ng1 - ng0).abs()
    assert bool((diff <= near).all()), "rank differs by more than the number of near-tied candidates: %s" % (
        torch.nonzero(diff > near).view(-1)[:8].tolist())
    exact = int((diff == 0).sum())
    assert exact >= 0.97 * len(diff), "only %d/%d ranks equal between tensor and fp32 paths" % (exact, len(diff))
    return ng1, ne1


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("E,H,Q", [(3000, 250, 300), (8000, 32, 700), (517, 30, 5), (256, 64, 128), (257, 16, 129)])
def test_tensor_path_matches_fp32_path(name, E, H, Q):
    """(8000, 32, 700): 6 x 32 = 192 tiles > 148 CTAs, so both accumulator stages and several barrier phases are
    exercised; (3000, 250, .): K = 250 / 500 is not a multiple of the 32-wide k-block; the last two sit on tile edges."""
    R = 7
    _, gm = _model(name, E, R, H, seed=E % 5)
    q = _queries(E, R, Q, seed=Q)
    indptr, idx = _filters(E, Q, seed=E + Q)
    for mode in ("head-batch", "tail-batch"):
        _compare(gm, q, mode, E)
        _compare(gm, q, mode, E, indptr, idx)


@pytest.mark.parametrize("name", NAMES)
def test_entity_chunks_and_ranges(name):
    """E above the 32768-row split-workspace chunk; sub-ranges accumulate to the full result."""
    from customknowledgegraphembedding_b200 import ops
    E, R, H, Q = 40000, 3, 16, 130
    _, gm = _model(name, E, R, H, seed=1)
    q = _queries(E, R, Q, seed=5)
    indptr, idx = _filters(E, Q, seed=9)
    ng, ne = _compare(gm, q, "tail-batch", E, indptr, idx)
    d = gm._desc()
    ts = torch.empty(Q, dtype=torch.float32, device="cuda")
    out = (ts, torch.zeros_like(ng), torch.zeros_like(ne))
    for lo, hi, acc in ((0, 255, False), (255, 33000, True), (33000, 33001, True), (33001, E, True)):
        ops.eval_rank(d, "tail-batch", gm.entity_embedding.detach(), gm.relation_embedding.detach(), None, q, lo, hi,
                      indptr, idx, out=out, accumulate=acc)
    assert torch.equal(out[1], ng) and torch.equal(out[2], ne)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_ranks(name):
    """Against the oracle's argsort ranks at a size where the tile grid has several query and entity tiles."""
    E, R, H = 1500, 5, 96
    om, gm = _model(name, E, R, H, seed=2)
    rng = np.random.RandomState(4)
    true = list({(int(h), int(r), int(t)) for h, r, t in zip(rng.randint(0, 60, 4000), rng.randint(0, R, 4000),
                                                             rng.randint(0, E, 4000))})
    test = true[:200]
    args = ko.Args(nentity=E, nrelation=R, test_batch_size=16)
    ref_metrics, ref_ranks, scores = ko.test_step(om, test, true, args, return_ranks=True, return_scores=True)
    got = type(gm).test_step(gm, test, true, args)
    ranks = gm.last_eval["ranks"]
    T = len(test)
    exact = 0
    for i in range(2 * T):
        h, r, t = test[i % T]
        s = scores[i].double().numpy()
        st = s[h if i < T else t]
        near = int((np.abs(s - st) <= 2e-5 * (12.0 + abs(st))).sum()) - 1
        assert abs(int(ranks[i]) - int(ref_ranks[i])) <= near
        exact += int(ranks[i] == ref_ranks[i])
    assert exact >= 0.97 * 2 * T
    if exact == 2 * T:
        for k in ref_metrics:
            assert got[k] == ref_metrics[k]


@pytest.mark.parametrize("name", NAMES)
def test_exact_ties(name):
    """Identical rows tie exactly with the true score through the tensor cores too (the true score is the
    diagonal of the same kind of tile): all-equal table -> n_greater 0, n_equal E-1; duplicated true row -> +1."""
    import customknowledgegraphembedding_b200 as kb
    E, R, H = 700, 2, 48
    de, dr, tr = ko.model_flags(name)
    m = kb.KGEModel(name, E, R, H, 12.0, de, dr, tr)
    with torch.no_grad():
        m.entity_embedding.fill_(0.25)
        m.relation_embedding.fill_(0.5)
    m = m.cuda()
    q = torch.tensor([[1, 0, 2], [3, 1, 4], [699, 1, 0]], dtype=torch.int64)
    for mode in ("head-batch", "tail-batch"):
        _, ng, ne = m.rank_queries(q, mode)
        assert ng.tolist() == [0, 0, 0] and ne.tolist() == [E - 1] * 3
    _, gm = _model(name, E, R, H, seed=3)
    with torch.no_grad():
        gm.entity_embedding[600] = gm.entity_embedding[17]
        gm.entity_embedding[333] = gm.entity_embedding[17]
    q = torch.tensor([[17, 0, 5], [5, 1, 17]], dtype=torch.int64)
    _, _, ne_h = gm.rank_queries(q[:1], "head-batch")
    _, _, ne_t = gm.rank_queries(q[1:], "tail-batch")
    assert ne_h.tolist() == [2] and ne_t.tolist() == [2]
    # ... and filtering one of the twins removes exactly it
    indptr = torch.tensor([0, 3], dtype=torch.int64)
    _, _, ne_f = gm.rank_queries(q[:1], "head-batch", indptr, torch.tensor([600, 17, 600], dtype=torch.int64))
    assert ne_f.tolist() == [1]


def test_filter_edge_cases_tensor_path():
    E, R, H = 300, 3, 16
    _, gm = _model("DistMult", E, R, H, seed=2)
    q = torch.tensor([[5, 1, 9]], dtype=torch.int64)
    _, ng, _ = gm.rank_queries(q, "tail-batch")
    s = _all_scores(gm, q.cuda(), "tail-batch", E)[0]
    better = torch.nonzero(s > s[9] + 1e-4).view(-1).cpu()
    assert int(ng[0]) >= len(better) > 0
    flt = torch.cat([better[:1], better[:1], torch.tensor([9]), better[:1], better[-1:]])
    _, ng2, _ = gm.rank_queries(q, "tail-batch", torch.tensor([0, len(flt)]), flt)
    assert int(ng2[0]) == int(ng[0]) - (2 if len(better) > 1 else 1)
    _, ng3, _ = gm.rank_queries(q, "tail-batch", torch.zeros(2, dtype=torch.int64), torch.zeros(0, dtype=torch.int64))
    assert int(ng3[0]) == int(ng[0])
