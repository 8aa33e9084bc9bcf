"""GPU: filtered ranking over a ROW-SHARDED entity table (kge_eval_rank_shard + RowShardedKGEModel) gives the
integers of the unsharded model.  World sizes 1..4 are emulated on one GPU by visiting the shards one after the
other (never several processes on one GPU); the real multi-process run is scratch/dist_check.py (NCCL)."""
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


def _csr(E, Q, seed):
    rng = np.random.RandomState(seed)
    lens = rng.randint(0, 9, Q)
    indptr = np.zeros(Q + 1, dtype=np.int64)
    np.cumsum(lens, out=indptr[1:])
    return torch.from_numpy(indptr).cuda(), torch.from_numpy(rng.randint(0, E, int(indptr[-1])).astype(np.int64)).cuda()


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
@pytest.mark.parametrize("world", [1, 2, 3])
def test_sharded_counts_equal_unsharded(name, world):
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H, Q = 1003, 5, 48, 150
    full = _model(name, E, R, H, seed=4)
    rng = np.random.RandomState(world)
    q = torch.from_numpy(np.stack([rng.randint(0, E, Q), rng.randint(0, R, Q), rng.randint(0, E, Q)], 1)).cuda()
    indptr, idx = _csr(E, Q, seed=7)
    shards = [kd.RowShardedKGEModel.from_full(full, rank_world=(r, world)) for r in range(world)]
    assert sum(s.entity_embedding.shape[0] for s in shards) == E
    for mode in ("head-batch", "tail-batch"):
        ts, ng, ne = full.rank_queries(q, mode, indptr, idx)
        # what gather_rows + the all-reduce assemble: the (head, tail) rows of every query
        qrows = full.entity_embedding.detach()[q[:, [0, 2]].reshape(-1)].view(Q, 2, -1).contiguous()
        tot_g, tot_e = torch.zeros_like(ng), torch.zeros_like(ne)
        for s in shards:
            # each shard's own contribution to the exchanged rows adds up to qrows exactly
            ts_s, g, e = s._rank_local(q, qrows, mode, indptr, idx)
            assert torch.equal(ts_s, ts), "true score must be bit-identical on every rank"
            tot_g += g
            tot_e += e
        assert torch.equal(tot_g, ng) and torch.equal(tot_e, ne)
    # the row exchange itself (world emulated: sum of the per-shard partial buffers)
    ids = q[:, [0, 2]].reshape(-1)
    parts = []
    for s in shards:
        mine = (ids >= s.lo) & (ids < s.hi)
        buf = torch.zeros((ids.numel(), full.entity_dim), device="cuda")
        buf[mine] = s.entity_embedding.detach()[ids[mine] - s.lo]
        parts.append(buf)
    assert torch.equal(sum(parts), full.entity_embedding.detach()[ids])


def test_shard_argument_checks():
    from customknowledgegraphembedding_b200 import dist as kd, ops
    full = _model("RotatE", 100, 3, 16, seed=1)
    s = kd.RowShardedKGEModel.from_full(full, rank_world=(1, 2))
    q = torch.tensor([[1, 0, 2]], dtype=torch.int64).cuda()
    qrows = full.entity_embedding.detach()[q[:, [0, 2]].reshape(-1)].view(1, 2, -1).contiguous()
    with pytest.raises(ValueError):  # candidate range outside the shard (KGE_ERR_INVALID_ARGUMENT)
        ops.eval_rank_shard(s._desc(), "tail-batch", s.entity_embedding.detach(), s.lo, s.relation_embedding.detach(),
                            None, q, qrows, 0, 100)
    with pytest.raises(NotImplementedError):
        s(q, "single")
