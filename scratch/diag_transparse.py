"""Diagnostic: TranSparse gradients at the bench width against the fp64 oracle, next to the fp32 oracle's own deviation.
|x| has a kink at 0: an entry of p_head * r - p_tail that is ~1e-9 can take either sign in fp32, and one flipped sign
moves a whole row of dX (and a column of dW) -- count how many rows / columns deviate and by how much."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import kge_oracle as ko
from customknowledgegraphembedding_b200.transparse import TranSparseModel

E, R, D, B, N = 5000, 11, 500, 64, 256
for tail in (False, True):
    torch.manual_seed(21)
    o = ko.OracleTranSparse(E, R, D, 12.0, project_tail=tail)
    m = TranSparseModel(E, R, D, 12.0, project_tail=tail)
    m.load_state_dict(o.state_dict())
    m = m.cuda()
    pos, neg, _, _ = ko.synthetic_batch(E, R, B, N, seed=22)
    o64 = ko.transparse_to_double(o)
    g = torch.Generator().manual_seed(6)
    ps, ns = m.both((pos.cuda(), neg.cuda()), "head-batch")
    up_p, up_n = torch.randn(ps.shape, generator=g), torch.randn(ns.shape, generator=g)
    (ps * up_p.cuda()).sum().add((ns * up_n.cuda()).sum()).backward()
    for model, dt in ((o, torch.float32), (o64, torch.float64)):
        ((model(pos, "single") * up_p.to(dt)).sum() + (model((pos, neg), "head-batch") * up_n.to(dt)).sum()).backward()
    for name in ("entity_embedding", "relation_embedding", "W"):
        want = getattr(o64, name).grad
        scale = float(want.abs().max())
        for who, got in (("gpu", getattr(m, name).grad.cpu().double()), ("fp32 oracle", getattr(o, name).grad.double())):
            err = (got - want).abs()
            bad = err > 2e-5 * scale
            rows = bad.reshape(-1, D).any(dim=1).sum().item()
            print("tail=%d %-18s %-11s scale %.3g max err %.3g  entries > 2e-5*scale: %d of %d (%d rows), median err %.2e"
                  % (tail, name, who, scale, float(err.max()), int(bad.sum()), err.numel(), rows, float(err.median())))
