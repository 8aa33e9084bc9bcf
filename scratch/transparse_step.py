"""Three TF train steps of TranSparse at the bench's sizes (bench.py TRANSPARSE), head-batch: the command profiled for
profiles/r2_transparse_launches.csv."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from customknowledgegraphembedding_b200.tf_compat import KerasAdam
from customknowledgegraphembedding_b200.transparse import TranSparseModel

c = bench.TRANSPARSE
torch.manual_seed(0)
m = TranSparseModel(c["nentity"], c["nrelation"], c["dim"], c["gamma"]).cuda()
opt = KerasAdam(learning_rate=c["lr"])
p, n, w, _ = bench.synthetic_batch(c["nentity"], c["nrelation"], c["batch"], c["negatives"], seed=9000)
p, n, w = p.cuda(), n.cuda(), w.cuda()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    loss = m.tf_train_step(opt, p, n, w, "head-batch")
    opt.iterations += 1
torch.cuda.synchronize()
print("loss", float(loss))
