"""Which positive / item / candidate row makes the warp-per-row sweep's dL/dq differ from the row-group sweep's?
(config 2, the batch of scratch/diag_grad.py)"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import kge_oracle as ko
import customknowledgegraphembedding_b200 as kb
from customknowledgegraphembedding_b200 import _lib, dist as kd

name, E, R, H, B, N, mode, gamma = "RotatE", 14541, 237, 1000, 1024, 256, "tail-batch", 9.0
om = ko.make_model(name, E, R, H, gamma, seed=0)
pos, neg, w, _ = ko.synthetic_batch(E, R, B, N, seed=5)
pos, neg, w = pos.cuda(), neg.cuda(), w.cuda()
lib = _lib.load()
gm = kb.KGEModel(name, E, R, H, gamma, True, False)
gm.load_state_dict(om.state_dict())
gm = gm.cuda()
opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=1e-3)
lo = kd.LossOpts(1, 0, 1, 0, 1.0, 0.0)
wsum = w.sum().reshape(1)
out = {}
for variant in (0, 1):
    lib.kge_train_set_sweep_variant(variant)
    st = kd.DataParallelStep(gm, opt)
    st._ensure(B, N, pos.device)
    st._sweep(mode, pos, neg, w, wsum, B, lo)
    torch.cuda.synchronize()
    out[variant] = {k: v.clone() for k, v in st._local.items()}
    if variant == 1:
        sws = st._sws.clone()
lib.kge_train_set_sweep_variant(1)
for k in ("S", "Q", "GH", "GT", "GR", "terms", "stats"):
    a, b = out[0][k].double(), out[1][k].double()
    d = (a - b).abs()
    print("%-6s max|diff| %.3e  max|ref| %.3e  argmax %d" % (k, d.max().item(), a.abs().max().item(), int(d.argmax())))
De = 2 * H
gh0, gh1 = out[0]["GH"].view(B, De).double(), out[1]["GH"].view(B, De).double()
rowerr = (gh0 - gh1).abs().amax(1) / gh0.abs().amax(1)
top = torch.argsort(rowerr, descending=True)[:8]
print("GH worst positives:", [(int(b), "%.2e" % rowerr[b].item()) for b in top])
b = int(top[0])
print("positive", b, pos[b].tolist(), "dups in neg[b]:", N - len(set(neg[b].tolist())),
      "neg contains h/t:", int((neg[b] == pos[b, 0]).sum()), int((neg[b] == pos[b, 2]).sum()))
# reference partials per item from the variant-1 Q (fp64)
parts, rpp = 8, 32
qdim = 2 * H
PA = sws[:B * parts * qdim * 4].view(torch.float32).view(B * parts, qdim).double()
off = (B * parts * qdim * 4 + 255) // 256 * 256
PM = sws[off:off + B * parts * 2 * 4].view(torch.float32).view(B * parts, 2).double()
Q = out[1]["Q"].view(B, qdim).double()
ent = gm.entity_embedding.detach().double()
with torch.no_grad():
    s_raw = gm((pos, neg), mode).double()  # K1 scores
for bb in [b, int(top[1])]:
    print("---- positive", bb)
    for p in range(parts):
        rows = neg[bb, p * rpp:(p + 1) * rpp]
        c = ent[rows]
        vr, vi = Q[bb, :H] - c[:, :H], Q[bb, H:] - c[:, H:]
        nrm = torch.sqrt(vr * vr + vi * vi)
        dr, di = vr / nrm, vi / nrm
        s = gamma - nrm.sum(1)
        a = s
        m_item = a.max()
        wgt = torch.sigmoid(s) * torch.exp(a - m_item)
        ref = torch.cat([-(wgt[:, None] * dr).sum(0), -(wgt[:, None] * di).sum(0)])
        got = PA[bb * parts + p]
        res = got - ref
        d_all = torch.cat([dr, di], 1)  # [32, 2H]
        proj = (d_all @ res) / (d_all * d_all).sum(1)   # weight error per row (rows are nearly orthogonal)
        j = int(proj.abs().argmax())
        print("item %d: m_item gpu %.6f ref %.6f | max|res| %.3e max|ref| %.3e | worst row n=%d id=%d weight err %.3e (weight %.3e) "
              "s_k1 %.5f s_ref %.5f | order of maxima: argmax a at n=%d" % (
                  p, PM[bb * parts + p, 0].item(), m_item.item(), res.abs().max().item(), ref.abs().max().item(),
                  p * rpp + j, int(rows[j]), proj[j].item(), wgt[j].item(), s_raw[bb, p * rpp + j].item(), s[j].item(),
                  p * rpp + int(a.argmax())))
