import sys, numpy as np, torch
sys.path.insert(0, '.')
from oracle import kge_oracle as ko
import customknowledgegraphembedding_b200 as kb
name, H, N, mode = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
E, R, B, lr = 300, 5, 12, 1e-3
gam = {"TransE": 12.0, "RotatE": 9.0, "InterHT": 24.0, "pRotatE": 9.0, "DistMult": 12.0, "ComplEx": 12.0}[name]
om = ko.make_model(name, E, R, H, gam, seed=11)
de, dr, tr = ko.model_flags(name)
gm = kb.KGEModel(name, E, R, H, gam, de, dr, tr); gm.load_state_dict(om.state_dict()); gm = gm.cuda()
args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=1.0, cuda=True)
o_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, om.parameters()), lr=lr)
g_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=lr)
pos, neg, w, _ = ko.synthetic_batch(E, R, B, N, seed=5)
want = ko.train_step(om, o_opt, iter([(pos, neg, w, mode)]), args)
got = kb.KGEModel.train_step(gm, g_opt, iter([(pos, neg, w, mode)]), args)
print(want, got)
for pn in ("entity_embedding", "relation_embedding"):
    g_ref = o_opt.state[getattr(om, pn)]["exp_avg"].numpy() * 10
    g_got = g_opt.state[getattr(gm, pn)]["exp_avg"].cpu().numpy() * 10
    err = np.abs(g_ref - g_got)
    tol = 1e-5 * (np.abs(g_ref) + np.abs(g_ref).max())
    bad = np.argwhere(err > tol)
    print(pn, "max|g|", np.abs(g_ref).max(), "bad", len(bad), "rows", sorted(set(bad[:, 0].tolist()))[:20])
    for r, c in bad[:8]:
        print("   ", r, c, g_ref[r, c], g_got[r, c])
    # relative error histogram for tiny gradients
    small = (np.abs(g_ref) < 1e-6) & (np.abs(g_ref) > 0)
    print("   tiny-gradient elements:", small.sum(), "exact zeros ref/got:", (g_ref == 0).sum(), (g_got == 0).sum())
print("pos", pos.tolist())
