"""GPU parity of K5 (filtered full-entity ranking by counting) against the oracle's upstream-style
test_step (materialise scores, add filter bias, argsort, locate the true entity).

Ranks are integers and must be equal.  The one legitimate source of difference is a candidate whose
fp32 score is within rounding distance of the true score: the two implementations sum the H terms in a
different order, so such a candidate can fall on either side.  The test therefore derives, from the
oracle's own scores, how many candidates of each query are that close (|s - s_true| <= AMBIG_TOL) and
requires |rank - rank_ref| <= that number -- and that it is 0 (bit-exact rank) for nearly every query."""
import numpy as np
import pytest
import torch

from oracle import kge_oracle as ko

pytestmark = pytest.mark.gpu

GAMMA = {"TransE": 12.0, "DistMult": 12.0, "ComplEx": 12.0, "RotatE": 9.0, "pRotatE": 9.0, "InterHT": 24.0}
AMBIG_TOL = 2e-5  # relative to (|gamma| + |score|)


def _pair(name, E, R, H, seed=0):
    import customknowledgegraphembedding_b200 as kb
    om = ko.make_model(name, E, R, H, GAMMA[name], seed=seed)
    de, dr, tr = ko.model_flags(name)
    gm = kb.KGEModel(name, E, R, H, GAMMA[name], de, dr, tr)
    gm.load_state_dict(om.state_dict())
    return om, gm.cuda()


def _graph(E, R, n_true, n_test, seed):
    rng = np.random.RandomState(seed)
    # few (h,r) / (r,t) keys so that filter lists are long
    heads = rng.randint(0, max(E // 8, 2), size=n_true)
    rels = rng.randint(0, R, size=n_true)
    tails = rng.randint(0, E, size=n_true)
    true = list({(int(h), int(r), int(t)) for h, r, t in zip(heads, rels, tails)})
    test = [true[i] for i in rng.choice(len(true), size=min(n_test, len(true)), replace=False)]
    return true, test


def _check_ranks(name, ranks, ref_ranks, scores, test, gamma):
    """scores: oracle [2*T, E] with filter bias already added."""
    T = len(test)
    exact = 0
    for i in range(2 * T):
        h, r, t = test[i % T]
        true_e = h if i < T else t
        s = scores[i].double().numpy()
        st = s[true_e]
        tol = AMBIG_TOL * (abs(gamma) + abs(st))
        near = int((np.abs(s - st) <= tol).sum()) - 1
        assert abs(int(ranks[i]) - int(ref_ranks[i])) <= near, (
            "%s query %d: rank %d vs reference %d with only %d candidates within rounding distance"
            % (name, i, ranks[i], ref_ranks[i], near))
        exact += int(ranks[i] == ref_ranks[i])
    assert exact >= 0.97 * 2 * T, "%s: only %d/%d ranks bit-exact" % (name, exact, 2 * T)
    return exact


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
@pytest.mark.parametrize("E,H", [(300, 32), (517, 100), (131, 30)])
def test_test_step_ranks_and_metrics(name, E, H):
    R = 5
    om, gm = _pair(name, E, R, H, seed=3)
    true, test = _graph(E, R, n_true=900, n_test=70, seed=E + H)
    args = ko.Args(nentity=E, nrelation=R, test_batch_size=16)
    ref_metrics, ref_ranks, scores = ko.test_step(om, test, true, args, return_ranks=True, return_scores=True)
    got_metrics = type(gm).test_step(gm, test, true, args)
    ranks = gm.last_eval["ranks"]
    assert ranks.shape == ref_ranks.shape
    exact = _check_ranks(name, ranks, ref_ranks, scores, test, GAMMA[name])
    assert set(got_metrics) == set(ref_metrics)
    if exact == len(ranks):  # identical ranks -> identical metrics, bit for bit (same summation order)
        for k in ref_metrics:
            assert got_metrics[k] == ref_metrics[k], k


def test_rank_counts_identities():
    """Structure of the counters, independent of any oracle."""
    from customknowledgegraphembedding_b200 import ops
    E, R, H = 1000, 4, 64
    _, gm = _pair("RotatE", E, R, H, seed=1)
    rng = np.random.RandomState(0)
    Q = 203
    queries = torch.from_numpy(np.stack([rng.randint(0, E, Q), rng.randint(0, R, Q), rng.randint(0, E, Q)], 1)).cuda()
    for mode in ("head-batch", "tail-batch"):
        ts, ng, ne = gm.rank_queries(queries, mode)
        # (a) splitting the entity range and accumulating gives the same counts
        out = (torch.empty_like(ts), torch.zeros_like(ng), torch.zeros_like(ne))
        d = gm._desc()
        for lo, hi, acc in ((0, 130, False), (130, 131, True), (131, E, True)):
            ops.eval_rank(d, mode, gm.entity_embedding.detach(), gm.relation_embedding.detach(), None, queries, lo, hi,
                          out=out, accumulate=acc)
        assert torch.equal(out[1], ng) and torch.equal(out[2], ne) and torch.equal(out[0], ts)
        # (b) 0 <= n_greater <= E-1, and a query order permutation permutes the result
        assert int(ng.min()) >= 0 and int(ng.max()) <= E - 1
        perm = torch.randperm(Q, device="cuda")
        ts2, ng2, ne2 = gm.rank_queries(queries[perm], mode)
        assert torch.equal(ng2, ng[perm]) and torch.equal(ts2, ts[perm])
        # (c) filtering every candidate that scores above the true one drives n_greater to 0
        with torch.no_grad():
            cand = torch.arange(E, device="cuda").repeat(Q, 1)
            s = gm((queries, cand), mode)  # K1 scores (ulp-different from K5's, so filter with a margin)
        col = 0 if mode == "head-batch" else 2
        s_true = s.gather(1, queries[:, col:col + 1])
        mask = s > (s_true - 1e-3)
        indptr = torch.zeros(Q + 1, dtype=torch.int64, device="cuda")
        indptr[1:] = mask.sum(1).cumsum(0)
        idx = cand[mask]
        _, ng3, ne3 = gm.rank_queries(queries, mode, indptr, idx)
        assert int(ng3.abs().max()) == 0 and int(ne3.abs().max()) == 0


def test_filter_edge_cases():
    """Duplicate filter entries, the true entity listed in its own filter, empty filters, Q = 1."""
    E, R, H = 150, 3, 16
    _, gm = _pair("TransE", E, R, H, seed=2)
    q = torch.tensor([[5, 1, 9]], dtype=torch.int64)
    ts, ng, ne = gm.rank_queries(q, "tail-batch")
    assert ng.shape == (1,)
    # entities that currently outrank the true tail
    with torch.no_grad():
        s = gm((q.cuda(), torch.arange(E).view(1, -1).cuda()), "tail-batch")[0]
    better = torch.nonzero(s > s[9] + 1e-4).view(-1).cpu()
    assert int(ng[0]) >= len(better) > 0
    flt = torch.cat([better[:1], better[:1], torch.tensor([9]), better[:1]])  # duplicate x3 + the true entity
    indptr = torch.tensor([0, len(flt)], dtype=torch.int64)
    _, ng2, _ = gm.rank_queries(q, "tail-batch", indptr, flt)
    assert int(ng2[0]) == int(ng[0]) - 1
    indptr0 = torch.zeros(2, dtype=torch.int64)
    _, ng3, _ = gm.rank_queries(q, "tail-batch", indptr0, torch.zeros(0, dtype=torch.int64))
    assert int(ng3[0]) == int(ng[0])


def test_exact_ties_are_reported():
    """All-equal embeddings: every candidate ties with the true one -> n_greater 0, n_equal E-1."""
    import customknowledgegraphembedding_b200 as kb
    E, R, H = 200, 2, 32
    m = kb.KGEModel("TransE", E, R, H, 12.0)
    with torch.no_grad():
        m.entity_embedding.fill_(0.25)
        m.relation_embedding.fill_(0.5)
    m = m.cuda()
    q = torch.tensor([[1, 0, 2], [3, 1, 4]], dtype=torch.int64)
    _, ng, ne = m.rank_queries(q, "head-batch")
    assert ng.tolist() == [0, 0] and ne.tolist() == [E - 1, E - 1]


def test_countries_auc_pr():
    """--countries branch of test_step: AUC-PR over the region candidates (scikit-learn), scores from K1."""
    from sklearn.metrics import average_precision_score
    E, R, H = 60, 2, 16
    om, gm = _pair("RotatE", E, R, H, seed=6)
    regions = [3, 11, 27, 40, 59]
    rng = np.random.RandomState(1)
    test = [(int(rng.randint(0, E)), int(rng.randint(0, R)), regions[int(rng.randint(0, 5))]) for _ in range(24)]
    args = ko.Args(nentity=E, nrelation=R)
    args.countries, args.regions = True, regions
    got = type(gm).test_step(gm, test, test, args)
    sample = torch.LongTensor([(h, r, c) for h, r, t in test for c in regions])
    y_true = np.array([1 if c == t else 0 for h, r, t in test for c in regions])
    with torch.no_grad():
        y_score = om(sample).squeeze(1).numpy()
    want = average_precision_score(y_true, y_score)
    assert set(got) == {"auc_pr"} and abs(got["auc_pr"] - want) < 1e-6


def test_both_modes_on_two_streams_equal_sequential():
    """rank_queries_both_modes (head-batch and tail-batch rankings side by side on two streams) gives the integers of
    the two sequential calls; test_step takes that path by default and `eval_two_streams = False` turns it off."""
    import customknowledgegraphembedding_b200 as kb
    E, R, H = 3000, 5, 200
    om, gm = _pair("RotatE", E, R, H, seed=4)
    true, test = _graph(E, R, n_true=4000, n_test=300, seed=9)
    q = torch.tensor(test, dtype=torch.int64)
    maps = gm._filter_maps(true)
    csr = {m: tuple(torch.from_numpy(a) for a in kb.KGEModel.build_filter_csr(maps, q.numpy(), m, E))
           for m in ("head-batch", "tail-batch")}
    for _ in range(3):
        both = gm.rank_queries_both_modes(q, csr)
        for mode in ("head-batch", "tail-batch"):
            want = gm.rank_queries(q, mode, *csr[mode])
            for a, b in zip(both[mode], want):
                assert torch.equal(a, b), mode
    args = ko.Args(nentity=E, nrelation=R, test_batch_size=16)
    m1 = type(gm).test_step(gm, test, true, args)
    r1 = gm.last_eval["ranks"].copy()
    gm.eval_two_streams = False
    m2 = type(gm).test_step(gm, test, true, args)
    assert np.array_equal(r1, gm.last_eval["ranks"]) and m1 == m2
