"""GPU: the ROW-SHARDED train step (RowShardedStep: compact-table sweep + kge_train_apply_shard) reproduces the
single-GPU step bit for bit.  Ranks are emulated on one GPU by running their pieces one after the other (the
NCCL run over real ranks is scratch/dist_check.py)."""
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


def _adam(m, lr):
    return torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=lr)


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
def test_world1_sharded_step_equals_plain_step(name):
    """One rank owning every row: exercises the unique / remap / compact-table sweep and the shard apply."""
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H, B, N, lr = 300, 5, 64, 16, 24, 1e-3
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=0.5, cuda=True)
    ref = _model(name, E, R, H, seed=2)
    ref_opt = _adam(ref, lr)
    sm = kd.RowShardedKGEModel.from_full(_model(name, E, R, H, seed=2), rank_world=(0, 1))
    step = kd.RowShardedStep(sm, _adam(sm, lr))
    for s in range(3):
        pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=s)
        batch = (pos.cuda(), neg.cuda(), w.cuda(), mode)
        want = kb.KGEModel.train_step(ref, ref_opt, iter([batch]), args)
        got = step.train_step(iter([batch]), args)
        assert got["loss"] == want["loss"], (s, got, want)
        assert torch.equal(sm.entity_embedding, ref.entity_embedding), s
        assert torch.equal(sm.relation_embedding, ref.relation_embedding), s


@pytest.mark.parametrize("E,B,N", [(300, 7, 13), (33, 4, 50), (1, 2, 3), (10_000_019, 64, 256), (4096, 16, 1000)])
def test_shard_compact_matches_torch_unique(E, B, N):
    """kge_shard_compact (bitmap mark + popcount scan) == torch.unique(sorted=True, return_inverse=True) of the ids a
    batch touches -- the sorted unique list, its length and the re-indexed batch; E not a multiple of 32, a table past
    2^23 words, a batch that touches every row (capacity = E)."""
    import ctypes as C
    from customknowledgegraphembedding_b200 import _lib, ops
    lib = _lib.load()
    g = torch.Generator().manual_seed(E + N)
    pos = torch.stack([torch.randint(0, E, (B,), generator=g), torch.randint(0, 5, (B,), generator=g),
                       torch.randint(0, E, (B,), generator=g)], 1).cuda()
    neg = torch.randint(0, E, (B, N), generator=g).cuda()
    if E > 1000:
        neg[0, 0], neg[0, 1], pos[1, 0] = E - 1, 0, E - 1   # the ends of the table
    cap = min(E, B * N + 2 * B)
    cpos, cneg = torch.empty_like(pos), torch.empty_like(neg)
    uniq = torch.full((cap,), -1, dtype=torch.int64, device="cuda")
    n = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.empty(lib.kge_shard_compact_workspace_bytes(E), dtype=torch.uint8, device="cuda")
    P = ops._ptr
    for _ in range(2):  # the workspace is reusable as is (the bitmap is cleared by the call)
        _lib.check(lib.kge_shard_compact(P(pos), P(neg), B, N, E, P(cpos), P(cneg), P(uniq), cap, P(n), P(ws), ws.numel(),
                                         ops._stream()), "kge_shard_compact")
    want, inv = torch.unique(torch.cat([pos[:, 0], pos[:, 2], neg.reshape(-1)]), sorted=True, return_inverse=True)
    assert int(n) == want.numel()
    assert torch.equal(uniq[:int(n)], want)
    assert torch.equal(cpos[:, 0], inv[:B]) and torch.equal(cpos[:, 2], inv[B:2 * B]) and torch.equal(cpos[:, 1], pos[:, 1])
    assert torch.equal(cneg.reshape(-1), inv[2 * B:])
    with pytest.raises(ValueError):
        _lib.check(lib.kge_shard_compact(P(pos), P(neg), B, N, E, P(cpos), P(cneg), P(uniq), cap - 1 if cap > 1 else 0, P(n),
                                         P(ws), ws.numel(), ops._stream()), "kge_shard_compact")


@pytest.mark.parametrize("De", [2000, 30, 4100, 8])
@pytest.mark.parametrize("world", [1, 3, 8])
def test_shard_fetch_rows_from_emulated_shards(De, world):
    """kge_shard_fetch_rows: rows of the global table pulled from `world` shards (here all on one GPU; over real ranks
    the pointers are symmetric-memory peer mappings) == table[ids]; ragged last shard, an empty shard, every row width
    path (16-byte lanes in 4 / 8 / 16 chunks, scalar)."""
    import ctypes as C
    from customknowledgegraphembedding_b200 import _lib, ops, dist as kd
    lib = _lib.load()
    E = 1001 if world < 8 else 9   # 9 rows over 8 ranks: shards of 2 rows, the last ranks empty
    torch.manual_seed(De + world)
    table = torch.randn(E, De, device="cuda")
    shards, ptrs = [], []
    for r in range(world):
        lo, hi = kd.shard_range(E, r, world)
        sh = table[lo:hi].clone() if hi > lo else torch.zeros(1, De, device="cuda")
        shards.append(sh)
        ptrs.append(C.c_void_p(sh.data_ptr()))
    ids = torch.unique(torch.randint(0, E, (700,), device="cuda"))
    cap = ids.numel() + 5
    uniq = torch.cat([ids, torch.full((5,), E + 7, dtype=torch.int64, device="cuda")])  # past n_unique: never read
    n = torch.tensor([ids.numel()], dtype=torch.int32, device="cuda")
    out = torch.full((cap, De), -7.0, device="cuda")
    arr = (C.c_void_p * world)(*ptrs)
    rows_per = (E + world - 1) // world
    _lib.check(lib.kge_shard_fetch_rows(ops._ptr(uniq), ops._ptr(n), cap, De, arr, world, rows_per, ops._ptr(out),
                                        ops._stream()), "kge_shard_fetch_rows")
    assert torch.equal(out[:ids.numel()], table[ids])
    assert bool((out[ids.numel():] == -7.0).all())


@pytest.mark.parametrize("name", ["RotatE", "TransE", "InterHT", "pRotatE"])
@pytest.mark.parametrize("world", [2, 3])
def test_emulated_ranks_kernel_path_equal_plain_step(name, world):
    """The same emulation with the kernel path of the sweep: touched ids by kge_shard_compact, rows pulled by
    kge_shard_fetch_rows from the (emulated) owners' shards instead of the torch.unique / all-to-all plumbing."""
    test_emulated_ranks_equal_plain_step(name, world, kernels=True)


@pytest.mark.parametrize("name", ["RotatE", "TransE", "InterHT", "pRotatE"])
@pytest.mark.parametrize("world", [2, 3])
def test_emulated_ranks_equal_plain_step(name, world, kernels=False):
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H, B, N, lr = 211, 5, 32, 12, 20, 1e-3
    half = B // world
    # uniform weights: coef = 1 / B_total on both sides, so the comparison is exact (with subsampling weights the
    # weight sum is reduced in a different order -> 1-ulp coefficients; covered with a tolerance by dist_check.py)
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=1.0, uni_weight=True, cuda=True)
    ref = _model(name, E, R, H, seed=5)
    ref_opt = _adam(ref, lr)
    full0 = _model(name, E, R, H, seed=5)
    shards = [kd.RowShardedKGEModel.from_full(full0, rank_world=(r, world)) for r in range(world)]
    steps = [kd.RowShardedStep(s, _adam(s, lr)) for s in shards]
    loss_opts = kd.LossOpts(1, 1, 1, 0, 1.0, 0.0)
    for it in range(3):
        pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=10 + it)
        pos, neg, w = pos.cuda(), neg.cuda(), w.cuda()
        want = kb.KGEModel.train_step(ref, ref_opt, iter([(pos, neg, w, mode)]), args)
        table = torch.cat([s.entity_embedding.detach() for s in shards])  # what the owners would serve
        wsum = w.sum().reshape(1)
        gathered = {}
        for r, st in enumerate(steps):
            sl = slice(r * half, (r + 1) * half)
            if kernels:
                import ctypes as C
                st.shard_ptrs = (C.c_void_p * world)(*[C.c_void_p(s.entity_embedding.data_ptr()) for s in shards])
            else:
                st.use_kernels = False
                st.fetch_rows = lambda ids, table=table: table[ids]  # the all-to-all, emulated
            st.world = 1
            st._shape = None
            st._ensure(half, N, pos.device)
            st._sweep(mode, pos[sl].contiguous(), neg[sl].contiguous(), w[sl].contiguous(), wsum, B, loss_opts)
            for k, v in st._local.items():
                gathered.setdefault(k, []).append(v.clone())
        gathered = {k: torch.cat(v) for k, v in gathered.items()}
        losses = []
        for r, st in enumerate(steps):
            st._gathered = {k: v.clone() for k, v in gathered.items()}  # the apply turns S into dL/ds in place
            m, opt = st.model, st.optimizer
            for p in (m.entity_embedding, m.relation_embedding) + ((m.modulus,) if name == "pRotatE" else ()):
                m._adam_state(opt, p)["step"] += 1
            adam = kd.AdamOpts(lr, 0.9, 0.999, 1e-8, it + 1, 0)
            out = torch.zeros(4, device="cuda")
            st._apply(mode, pos, neg, loss_opts, adam, out)
            losses.append(out[2].item())
        assert all(l == want["loss"] for l in losses), (losses, want)
        now = torch.cat([s.entity_embedding.detach() for s in shards])
        assert torch.equal(now, ref.entity_embedding.detach()), it
        for s in shards:
            assert torch.equal(s.relation_embedding, ref.relation_embedding)
            if name == "pRotatE":
                assert torch.equal(s.modulus, ref.modulus)


@pytest.mark.parametrize("world", [1, 2, 3])
def test_peer_exchange_places_every_row_where_it_is_needed(world):
    """kge_train_exchange with the ranks emulated on one GPU (every "peer" block is a plain tensor here; over real
    ranks the same pointers come from symmetric memory -- scratch/dist_check.py): ids, S, Q, GR and the scalars land
    in every rank's gathered buffers at slot rank * B + b; GH / GT rows only at the owner of the head / tail."""
    import ctypes as C
    from customknowledgegraphembedding_b200 import _lib, ops
    from customknowledgegraphembedding_b200._lib import StepBuffers
    E, R, H, B, N = 50, 3, 8, 5, 7
    m = _model("RotatE", E, R, H, seed=1)
    sizes = {"S": N, "Q": 2 * H, "GH": 2 * H, "GT": 2 * H, "GR": H, "GM": 1, "stats": 2, "terms": 2}
    names = ("S", "Q", "GH", "GT", "GR", "GM", "stats", "terms")
    g = torch.Generator().manual_seed(3)
    local = [{k: torch.randn(B * s, generator=g).cuda() for k, s in sizes.items()} for _ in range(world)]
    pos = [torch.stack([torch.randint(0, E, (B,), generator=g), torch.randint(0, R, (B,), generator=g),
                        torch.randint(0, E, (B,), generator=g)], 1).cuda() for _ in range(world)]
    neg = [torch.randint(0, E, (B, N), generator=g).cuda() for _ in range(world)]
    SENT = -777.0
    gathered = [{k: torch.full((world * B * s,), SENT, device="cuda") for k, s in sizes.items()} for _ in range(world)]
    gpos = [torch.full((world * B, 3), -1, dtype=torch.int64, device="cuda") for _ in range(world)]
    gneg = [torch.full((world * B, N), -1, dtype=torch.int64, device="cuda") for _ in range(world)]
    tables = (StepBuffers * world)(*[StepBuffers(*[C.c_void_p(gathered[p][k].data_ptr()) for k in names]) for p in range(world)])
    ppos = (C.c_void_p * world)(*[C.c_void_p(t.data_ptr()) for t in gpos])
    pneg = (C.c_void_p * world)(*[C.c_void_p(t.data_ptr()) for t in gneg])
    lib = _lib.load()
    for r in range(world):
        loc = StepBuffers(*[C.c_void_p(local[r][k].data_ptr()) for k in names])
        _lib.check(lib.kge_train_exchange(C.byref(m._desc()), ops._ptr(pos[r]), ops._ptr(neg[r]), B, N, C.byref(loc), r, world,
                                          tables, ppos, pneg, ops._stream()), "kge_train_exchange")
    torch.cuda.synchronize()
    rows = (E + world - 1) // world
    for p in range(world):
        assert torch.equal(gpos[p], torch.cat(pos)) and torch.equal(gneg[p], torch.cat(neg))
        for k in ("S", "Q", "GR", "GM", "stats", "terms"):
            assert torch.equal(gathered[p][k], torch.cat([local[r][k] for r in range(world)])), (p, k)
        for k, col in (("GH", 0), ("GT", 2)):
            got = gathered[p][k].view(world * B, -1)
            want = torch.cat([local[r][k] for r in range(world)]).view(world * B, -1)
            owner = torch.cat(pos)[:, col] // rows
            mine = owner == p
            assert torch.equal(got[mine], want[mine]), (p, k)
            assert bool((got[~mine] == SENT).all()), (p, k)  # rows nobody here will read were not sent
    with pytest.raises(ValueError):
        lib_rc = lib.kge_train_exchange(C.byref(m._desc()), ops._ptr(pos[0]), ops._ptr(neg[0]), B, N, C.byref(loc), world, world,
                                        tables, ppos, pneg, ops._stream())
        _lib.check(lib_rc, "kge_train_exchange")


def test_updated_rows_are_stored_into_the_replicas():
    """kge_train_apply_peers: every updated entity row is also written to the peer tables (two extra buffers on the
    same GPU stand in for the other ranks' replicas); rows outside the applied range stay as they were."""
    import ctypes as C
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H, B, N, lr = 90, 4, 16, 8, 11, 1e-3
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=1.0, cuda=True)
    m = _model("RotatE", E, R, H, seed=4)
    before = m.entity_embedding.detach().clone()
    replicas = [before.clone(), before.clone()]
    lo, hi = 10, 70

    class Peers(kd.DataParallelStep):
        def _apply(self, mode, pos_all, neg_all, loss_opts, adam, out, **kw):
            ptrs = (C.c_void_p * 2)(*[C.c_void_p(r.data_ptr()) for r in replicas])
            super()._apply(mode, pos_all, neg_all, loss_opts, adam, out, row_peers=ptrs)

    step = Peers(m, _adam(m, lr))
    step.lo, step.hi = lo, hi  # the rows this "rank" owns
    pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=8)
    step.train_step(iter([(pos.cuda(), neg.cuda(), w.cuda(), mode)]), args)
    now = m.entity_embedding.detach()
    assert not torch.equal(now[lo:hi], before[lo:hi])
    for r in replicas:
        assert torch.equal(r[lo:hi], now[lo:hi])
        assert torch.equal(r[:lo], before[:lo]) and torch.equal(r[hi:], before[hi:])
    assert torch.equal(now[:lo], before[:lo]) and torch.equal(now[hi:], before[hi:])
