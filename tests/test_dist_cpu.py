"""CPU, world_size 2 over gloo: the host logic of customknowledgegraphembedding_b200.dist (shard ranges,
all-reduce of the weight sum and of the rank counters, all-gather of the per-positive buffers and batches,
owner-parallel apply, broadcast of the owned rows) with oracle stand-ins for the kernel calls."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_covers_everything():
    from customknowledgegraphembedding_b200.dist import shard_range
    for n in (1, 7, 8, 14541, 10_000_000):
        for world in (1, 2, 3, 8):
            prev = 0
            for r in range(world):
                lo, hi = shard_range(n, r, world)
                assert lo == prev and hi >= lo
                prev = hi
            assert prev == n


def test_row_sharded_init_does_not_depend_on_the_sharding():
    """Block-wise streams: the concatenation of the shards is the same table for every world size."""
    from customknowledgegraphembedding_b200 import dist as kd

    class M(kd.RowShardedKGEModel):
        INIT_BLOCK = 7
    full = M("RotatE", 50, 3, 4, 9.0, True, False, rank_world=(0, 1), seed=2)
    for world in (2, 3, 8):
        parts = [M("RotatE", 50, 3, 4, 9.0, True, False, rank_world=(r, world), seed=2) for r in range(world)]
        assert torch.equal(torch.cat([p.entity_embedding for p in parts]), full.entity_embedding)
        assert all(torch.equal(p.relation_embedding, full.relation_embedding) for p in parts)
        assert all(p.nentity == 50 for p in parts)


def _oracle_counts(om, queries, mode, indptr, idx, entity_lo, entity_hi):
    """CPU stand-in for kge_eval_rank over [entity_lo, entity_hi): same contract as KGEModel.rank_queries."""
    Q, E = queries.shape[0], om.nentity
    cand = torch.arange(E).repeat(Q, 1)
    with torch.no_grad():
        s = om((queries, cand), mode)
    col = 0 if mode == "head-batch" else 2
    true_e = queries[:, col]
    ts = s.gather(1, true_e.view(-1, 1)).view(-1)
    keep = torch.zeros(Q, E, dtype=torch.bool)
    keep[:, entity_lo:entity_hi] = True
    keep[torch.arange(Q), true_e] = False
    if indptr is not None:
        for q in range(Q):
            keep[q, idx[indptr[q]:indptr[q + 1]]] = False
    ng = ((s > ts.view(-1, 1)) & keep).sum(1)
    ne = ((s == ts.view(-1, 1)) & keep).sum(1)
    return ts, ng, ne


def _worker(rank, world, port, tmp):
    sys.path.insert(0, REPO)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    from oracle import kge_oracle as ko
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import dist as kd

    E, R, H, B, N = 61, 4, 8, 6, 5
    om = ko.make_model("RotatE", E, R, H, 9.0, seed=3)
    gm = kb.KGEModel("RotatE", E, R, H, 9.0, True, False)   # stays on the CPU: kernels are stood in
    gm.load_state_dict(om.state_dict())

    # ---------------- eval: sharded counts + all-reduce == unsharded ----------------
    rng = np.random.RandomState(0)
    true = [(int(a), int(b), int(c)) for a, b, c in zip(rng.randint(0, 9, 80), rng.randint(0, R, 80), rng.randint(0, E, 80))]
    test = true[:11]
    args = ko.Args(nentity=E, nrelation=R)
    local = lambda q, mode, ip, ix, entity_lo, entity_hi: _oracle_counts(om, q, mode, ip, ix, entity_lo, entity_hi)
    got = kd.test_step(gm, test, true, args, local_fn=local)
    want, ranks = ko.test_step(om, test, true, args, return_ranks=True)
    assert np.array_equal(gm.last_eval["ranks"], ranks), (gm.last_eval["ranks"], ranks)
    assert got == want

    # ---------------- eval over a ROW-SHARDED table: row exchange + sharded counts == unsharded ----------------
    sm = kd.RowShardedKGEModel.from_full(gm)
    assert (sm.lo, sm.hi) == kd.shard_range(E, rank, world) and sm.entity_embedding.shape[0] == sm.hi - sm.lo
    ids = torch.tensor([0, E - 1, 30, 31, 5, 5])
    assert torch.equal(sm.gather_rows(ids), om.entity_embedding.detach()[ids])   # owners + zeros, summed: exact

    def shard_local(q, qrows, mode, ip, ix):
        want_rows = om.entity_embedding.detach()[q[:, [0, 2]].reshape(-1)].view(q.shape[0], 2, -1)
        assert torch.equal(qrows, want_rows)
        return _oracle_counts(om, q, mode, ip, ix, sm.lo, sm.hi)
    got_s = kb.KGEModel.test_step(sm, test, true, args,
                                  rank_fn=lambda q, mode, ip, ix: sm.rank_queries(q, mode, ip, ix, local_fn=shard_local))
    assert np.array_equal(sm.last_eval["ranks"], ranks) and got_s == want

    # ---------------- row-sharded train step: the id / row all-to-all (each rank asks for different rows) ----------
    rstep = kd.RowShardedStep(sm, torch.optim.Adam(filter(lambda p: p.requires_grad, sm.parameters()), lr=1e-2))
    want_ids = torch.unique(torch.from_numpy(np.random.RandomState(50 + rank).randint(0, E, 40 + 7 * rank)))
    assert torch.equal(rstep.fetch_rows(want_ids), om.entity_embedding.detach()[want_ids])
    assert rstep.last_exchange["rows_requested"] == want_ids.numel()
    only_low = torch.arange(0, 5)  # every rank asks rank 0 only; rank 1 serves nothing
    assert torch.equal(rstep.fetch_rows(only_low), om.entity_embedding.detach()[only_low])
    assert torch.equal(rstep.fetch_rows(torch.zeros(0, dtype=torch.int64)), torch.zeros(0, sm.entity_dim))

    # ---------------- train: sweep local / gather / apply owned rows / broadcast ----------------
    class Step(kd.DataParallelStep):
        def _sweep(self, mode, pos, neg, w, wsum, B_total, loss_opts):
            self.seen_wsum = float(wsum)           # must be the GLOBAL sum
            self._local["terms"][:] = float(self.rank)  # marker: gathered in rank order

        def _apply(self, mode, pos_all, neg_all, loss_opts, adam, out):
            # oracle gradient of the GLOBAL batch, Adam on the owned entity rows only (+ all relation rows)
            w_all = torch.cat(self.w_all)
            ref = ko.make_model("RotatE", E, R, H, 9.0, seed=3)
            ref.load_state_dict(self.model.state_dict())
            neg_s = ref((pos_all, neg_all), mode)
            pos_s = ref(pos_all)
            _, _, loss = ko.kge_loss(neg_s, pos_s, w_all, negative_adversarial_sampling=True)
            loss.backward()
            lr = adam.lr
            with torch.no_grad():
                g = ref.entity_embedding.grad
                self.model.entity_embedding[self.lo:self.hi] -= lr * torch.sign(g[self.lo:self.hi])  # Adam step 1
                self.model.relation_embedding -= lr * torch.sign(ref.relation_embedding.grad)
            out[2] = loss.detach()
            self.terms_seen = self._gathered["terms"].clone()

    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=1e-2)
    step = Step(gm, opt)
    assert (step.lo, step.hi) == kd.shard_range(E, rank, world)
    batches = [ko.synthetic_batch(E, R, B, N, seed=10 + r, mode="tail-batch") for r in range(world)]
    step.w_all = [b[2] for b in batches]
    targs = ko.Args(negative_adversarial_sampling=True)
    log = step.train_step(iter([batches[rank]]), targs)
    assert abs(step.seen_wsum - float(sum(b[2].sum() for b in batches))) < 1e-5
    # gathered in rank order, B rows of 2 terms each
    assert torch.equal(step.terms_seen, torch.arange(world).repeat_interleave(2 * B).float())
    # every rank ends with the same, fully updated table
    full = [torch.empty_like(gm.entity_embedding.data) for _ in range(world)]
    dist.all_gather(full, gm.entity_embedding.data.contiguous())
    assert all(torch.equal(full[0], f) for f in full)
    moved = (gm.entity_embedding.data - om.entity_embedding.data).abs().amax(1) > 0
    pos_all = torch.cat([b[0] for b in batches]); neg_all = torch.cat([b[1] for b in batches])
    touched = torch.zeros(E, dtype=torch.bool)
    touched[pos_all[:, 0]] = True; touched[pos_all[:, 2]] = True; touched[neg_all.reshape(-1)] = True
    assert torch.equal(moved, touched)   # rows of BOTH shards moved, on both ranks
    # the loss is the single-process loss of the concatenated batch
    ref = ko.make_model("RotatE", E, R, H, 9.0, seed=3)
    _, _, want_loss = ko.kge_loss(ref((pos_all, neg_all), "tail-batch"), ref(pos_all), torch.cat(step.w_all),
                                  negative_adversarial_sampling=True)
    assert abs(log["loss"] - want_loss.item()) < 1e-6

    # ---------------- unequal batch shapes across ranks are refused, on every rank, before any exchange -------------
    step2 = Step(gm, opt)
    step2.w_all = step.w_all
    bad = ko.synthetic_batch(E, R, B + rank, N, seed=99, mode="tail-batch")
    try:
        step2.train_step(iter([bad]), targs)
        raise AssertionError("unequal B was accepted")
    except ValueError as e:
        assert "different batch shapes" in str(e), e
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(tmp, "ok%d" % rank), "w").close()


def test_two_ranks_gloo(tmp_path):
    world = 2
    port = 29500 + (os.getpid() % 1000)
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(os.path.join(str(tmp_path), "ok%d" % r)) for r in range(world))
