"""N1, the on-GPU producer of the train batch (customknowledgegraphembedding_b200/sampler.py, csrc/kge_sampler.cu)
against the oracle's restatement of the upstream TrainDataset (U11) and of the kernel's Philox draw scheme."""
import numpy as np
import pytest
import torch

from oracle import kge_oracle as ko


def _triples(E, R, T, seed):
    rng = np.random.RandomState(seed)
    # few (h, r) / (r, t) keys so that true lists are long; duplicates on purpose
    tr = np.stack([rng.randint(0, max(E // 6, 2), T), rng.randint(0, R, T), rng.randint(0, E, T)], 1)
    tr = np.concatenate([tr, tr[:7]])
    return [tuple(int(x) for x in t) for t in tr]


def test_philox_known_answers():
    """Random123 kat_vectors, philox4x32 with 10 rounds."""
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
             (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for c, k, want in kats:
        assert ko.philox4x32_10(c, k) == want


def test_index_matches_upstream_dataset_structures():
    """Weights bit-equal to torch.sqrt(1 / torch.Tensor([count...])); true sets equal to true_head / true_tail."""
    from customknowledgegraphembedding_b200.sampler import TrainTripleIndex
    E, R = 60, 4
    tr = _triples(E, R, 300, seed=1)
    idx = TrainTripleIndex(tr, E, R, device="cpu")
    count = ko.count_frequency(tr)
    want_w = torch.cat([ko.subsampling_weight_of(t, count) for t in tr])
    assert torch.equal(idx.weights, want_w)
    true_head, true_tail = ko.get_true_head_and_tail(tr)
    comp_h = sorted((r * E + t) * E + int(h) for (r, t), hs in true_head.items() for h in hs)
    comp_t = sorted((r * E + h) * E + int(t) for (h, r), ts in true_tail.items() for t in ts)
    assert idx.true_sorted["head-batch"].tolist() == comp_h
    assert idx.true_sorted["tail-batch"].tolist() == comp_t
    with pytest.raises(ValueError):
        TrainTripleIndex([(0, 0, E)], E, R, device="cpu")


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["head-batch", "tail-batch"])
def test_gpu_sampler_equals_reference_draws(mode):
    from customknowledgegraphembedding_b200.sampler import TrainTripleIndex, sample_negatives
    E, R, N = 50, 3, 37
    tr = _triples(E, R, 400, seed=2)       # dense: ~1/6 of the candidates of a key are known
    idx = TrainTripleIndex(tr, E, R)
    pos = torch.tensor(tr[:23], dtype=torch.int64).cuda()
    got = sample_negatives(idx, pos, N, mode, seed=0x1234567890ABCDEF, step=5, first_index=(1 << 33) + 11)
    want = ko.sample_negatives_reference(tr[:23], N, mode, E, tr, 0x1234567890ABCDEF, 5, (1 << 33) + 11)
    assert np.array_equal(got.cpu().numpy(), want)
    # no negative forms a known triple
    true = set(tr)
    for (h, r, t), row in zip(tr[:23], got.tolist()):
        assert all(((e, r, t) if mode == "head-batch" else (h, r, e)) not in true for e in row)
    # a batch split over ranks is the same batch
    a = sample_negatives(idx, pos[:10], N, mode, 0x1234567890ABCDEF, 5, first_index=(1 << 33) + 11)
    b = sample_negatives(idx, pos[10:], N, mode, 0x1234567890ABCDEF, 5, first_index=(1 << 33) + 21)
    assert torch.equal(torch.cat([a, b]), got)
    with pytest.raises(ValueError):
        sample_negatives(idx, pos, N, "single", 0, 0)


@pytest.mark.gpu
def test_gpu_sampler_is_uniform_over_the_allowed_entities():
    from customknowledgegraphembedding_b200.sampler import TrainTripleIndex, sample_negatives
    E, R = 40, 2
    tr = [(0, 0, t) for t in range(0, 40, 3)] + [(1, 1, 5)]
    idx = TrainTripleIndex(tr, E, R)
    pos = torch.tensor([tr[0]] * 64, dtype=torch.int64).cuda()
    neg = sample_negatives(idx, pos, 4096, "tail-batch", seed=7, step=0)
    counts = torch.bincount(neg.reshape(-1), minlength=E).cpu().numpy().astype(np.float64)
    known = np.zeros(E, dtype=bool)
    known[list(range(0, 40, 3))] = True
    assert counts[known].sum() == 0
    allowed = counts[~known]
    expect = allowed.sum() / allowed.size
    chi2 = ((allowed - expect) ** 2 / expect).sum()
    assert chi2 < 60.0, chi2     # 25 degrees of freedom: P(chi2 > 60) ~ 1e-4


@pytest.mark.gpu
def test_gpu_iterator_feeds_train_step():
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200.sampler import TrainTripleIndex, BidirectionalGpuIterator
    E, R, H, B, N = 80, 4, 32, 16, 24
    tr = _triples(E, R, 100, seed=3)
    idx = TrainTripleIndex(tr, E, R)
    it = BidirectionalGpuIterator(idx, N, B, seed=1)
    model = kb.KGEModel("RotatE", E, R, H, 9.0, True, False).cuda()
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, model.parameters()), lr=1e-3)
    args = ko.Args(negative_adversarial_sampling=True, cuda=True)
    seen, modes, losses = [], [], []
    for step in range(16):   # 107 triples / 16 per batch: two epochs per mode
        pos, neg, w, mode = next(it)
        assert pos.is_cuda and neg.shape == (pos.shape[0], N) and w.shape == (pos.shape[0],)
        modes.append(mode)
        if mode == "head-batch":
            seen.append(pos.cpu())
        losses.append(kb.KGEModel.train_step(model, opt, iter([(pos, neg, w, mode)]), args)["loss"])
    assert modes[:4] == ["head-batch", "tail-batch", "head-batch", "tail-batch"]
    first_epoch = torch.cat(seen[:7])   # 7 head batches cover the 107 triples once
    assert sorted(map(tuple, first_epoch.tolist())) == sorted(tr)
    assert all(np.isfinite(losses))
    # two ranks see disjoint shares whose union is the single-process batch
    full = BidirectionalGpuIterator(idx, N, B, seed=5)
    r0 = BidirectionalGpuIterator(idx, N, B // 2, seed=5, rank=0, world=2)
    r1 = BidirectionalGpuIterator(idx, N, B // 2, seed=5, rank=1, world=2)
    for _ in range(3):
        p, n, w, m = next(full)
        p0, n0, w0, m0 = next(r0)
        p1, n1, w1, m1 = next(r1)
        assert m == m0 == m1
        assert torch.equal(torch.cat([p0, p1]), p) and torch.equal(torch.cat([n0, n1]), n) and torch.equal(torch.cat([w0, w1]), w)


def test_iterator_shares_are_equal_across_ranks_over_epochs():
    """world > 1: drop_last is forced, so every rank gets exactly B positives in every step -- also across the epoch
    boundary where a short last batch used to be split 890/889 (or leave a rank empty).  Host logic only (CPU index)."""
    from customknowledgegraphembedding_b200.sampler import TrainTripleIndex, BidirectionalGpuIterator
    E, R, B, world = 50, 3, 8, 3
    tr = _triples(E, R, 70, seed=4)          # 70+ triples, global batch 24: 2 full batches per epoch, remainder dropped
    idx = TrainTripleIndex(tr, E, R, device="cpu")
    its = [BidirectionalGpuIterator(idx, 4, B, seed=2, rank=r, world=world) for r in range(world)]
    assert all(it.drop_last for it in its)
    for step in range(9):                     # three epochs per mode
        mode = "head-batch" if step % 2 == 0 else "tail-batch"
        shares = [it._next_positions(mode) for it in its]
        assert [len(s[0]) for s in shares] == [B] * world
        assert [s[1] for s in shares] == [shares[0][1] + r * B for r in range(world)]   # consecutive global indices
        assert len(set(torch.cat([s[0] for s in shares]).tolist())) == B * world        # disjoint
    with pytest.raises(ValueError):
        BidirectionalGpuIterator(idx, 4, B, rank=0, world=2, drop_last=False)
    with pytest.raises(ValueError):
        BidirectionalGpuIterator(idx, 4, 64, rank=0, world=2)
