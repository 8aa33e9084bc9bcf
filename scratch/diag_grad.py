"""Where does the config-2 gradient differ from the fp64 twin?  GPU fused (sweep variants 1 / 0), GPU v1 kernels, fp32 oracle."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import kge_oracle as ko
import customknowledgegraphembedding_b200 as kb
from customknowledgegraphembedding_b200 import _lib

name, E, R, H, B, N, mode, gamma = "RotatE", 14541, 237, 1000, int(os.environ.get("DB", 1024)), 256, os.environ.get("DMODE", "tail-batch"), 9.0
seed = int(os.environ.get("DSEED", 0))
om = ko.make_model(name, E, R, H, gamma, seed=seed)
o64 = ko.to_double(om)
pos, neg, w, _ = ko.synthetic_batch(E, R, B, N, seed=seed + 5)
args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=1.0, cuda=True)
lib = _lib.load()

def gpu_grad(variant, det):
    gm = kb.KGEModel(name, E, R, H, gamma, True, False)
    gm.load_state_dict(om.state_dict())
    gm = gm.cuda()
    gm.deterministic = det
    lib.kge_train_set_sweep_variant(variant)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, gm.parameters()), lr=1e-3)
    log = kb.KGEModel.train_step(gm, opt, iter([(pos, neg, w, mode)]), args)
    lib.kge_train_set_sweep_variant(1)
    return opt.state[gm.entity_embedding]["exp_avg"].cpu().double().numpy() * 10.0, log["loss"]

res = {}
res["gpu_fused_warp"] = gpu_grad(1, 0)
res["gpu_fused_rowgroup"] = gpu_grad(0, 0)
res["gpu_v1_atomic"] = gpu_grad(1, 2)
res["gpu_v1_det"] = gpu_grad(1, 3)
t = time.time()
o_opt = torch.optim.Adam(filter(lambda p: p.requires_grad, om.parameters()), lr=1e-3)
l32 = ko.train_step(om, o_opt, iter([(pos, neg, w, mode)]), args)["loss"]
res["oracle_fp32"] = (o_opt.state[om.entity_embedding]["exp_avg"].double().numpy() * 10.0, l32)
print("fp32 oracle %.1fs" % (time.time() - t), flush=True)
t = time.time()
opt64 = torch.optim.Adam(filter(lambda p: p.requires_grad, o64.parameters()), lr=1e-3)
l64 = ko.train_step(o64, opt64, iter([(pos, neg, w.double(), mode)]), args)["loss"]
g64 = opt64.state[o64.entity_embedding]["exp_avg"].numpy() * 10.0
print("fp64 oracle %.1fs loss %.9f" % (time.time() - t, l64), flush=True)
gmax = np.abs(g64).max()
uses = np.bincount(neg.numpy().ravel(), minlength=E)
for k, (g, loss) in res.items():
    err = np.abs(g - g64) / gmax
    flat = np.argsort(err.ravel())[::-1][:6]
    print("%-20s loss %.9f  max err/gmax %.3e  p99.9999 %.3e  mean %.3e  >5e-5: %d  >1e-5: %d" % (
        k, loss, err.max(), np.quantile(err, 0.999999), err.mean(), (err > 5e-5).sum(), (err > 1e-5).sum()))
    for f in flat:
        e, c = divmod(int(f), 2 * H)
        print("      row %5d col %4d err %.3e g64 %.3e got %.3e uses %d" % (e, c, err[e, c], g64[e, c] / gmax, g[e, c] / gmax, uses[e]))
# decompose the worst element of the fused warp variant
g = res["gpu_fused_warp"][0]
err = np.abs(g - g64) / gmax
e, c = divmod(int(err.argmax()), 2 * H)
k = c % H
print("worst element row %d col %d (coordinate %d, plane %d)" % (e, c, k, c // H))
bs, ns = np.nonzero(neg.numpy() == e)
E64, R64 = o64.entity_embedding.detach(), o64.relation_embedding.detach()
rng = o64.embedding_range.item() / 3.14159265358979323846
for b, n in zip(bs, ns):
    h, r, t = pos[b].tolist()
    fixed = E64[h] if mode == "tail-batch" else E64[t]
    ph = R64[r, k] / rng
    cr, sr = torch.cos(ph), torch.sin(ph)
    if mode == "tail-batch":
        qr, qi = fixed[k] * cr - fixed[H + k] * sr, fixed[k] * sr + fixed[H + k] * cr
    else:
        qr, qi = cr * fixed[k] + sr * fixed[H + k], cr * fixed[H + k] - sr * fixed[k]
    vr, vi = qr - E64[e, k], qi - E64[e, H + k]
    print("   use b=%d n=%d |v|=%.3e vr=%.3e vi=%.3e" % (b, n, float(torch.sqrt(vr * vr + vi * vi)), float(vr), float(vi)))
print("positives with this entity as head/tail:", int((pos[:, 0] == e).sum()), int((pos[:, 2] == e).sum()))
