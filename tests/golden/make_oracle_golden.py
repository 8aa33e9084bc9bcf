"""Write ``oracle_models.npz``: oracle outputs for the five models whose reference source is absent
(SURVEY 8(a) U3-U7).  These are **drift detectors, not reference pins** ("parity unpinned"): they
record what oracle/kge_oracle.py produced when it was written, so a later edit of the oracle or a
torch upgrade that changes results is noticed.

    python tests/golden/make_oracle_golden.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import kge_oracle as ko  # noqa: E402

E, R, H, B, N = 61, 5, 20, 4, 7
GAMMA = {"TransE": 12.0, "DistMult": 12.0, "ComplEx": 12.0, "RotatE": 9.0, "pRotatE": 9.0, "InterHT": 24.0}


def main():
    torch.set_num_threads(1)
    out = {}
    for name in ko.MODEL_NAMES:
        m = ko.make_model(name, E, R, H, GAMMA[name], seed=11)
        pos, neg, w, _ = ko.synthetic_batch(E, R, B, N, seed=5)
        out[name + "/entity_embedding"] = m.entity_embedding.detach().numpy().copy()
        out[name + "/relation_embedding"] = m.relation_embedding.detach().numpy().copy()
        out[name + "/positive_sample"] = pos.numpy()
        out[name + "/negative_sample"] = neg.numpy()
        out[name + "/subsampling_weight"] = w.numpy()
        with torch.no_grad():
            out[name + "/single"] = m(pos, "single").numpy()
            out[name + "/head-batch"] = m((pos, neg), "head-batch").numpy()
            out[name + "/tail-batch"] = m((pos, neg), "tail-batch").numpy()
        # one upstream-style train step (adv, alpha .5, lr 1e-2) -> loss and updated tables
        opt = torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=1e-2)
        args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=0.5)
        log = ko.train_step(m, opt, iter([(pos, neg, w, "tail-batch")]), args)
        out[name + "/loss"] = np.float32(log["loss"])
        out[name + "/entity_embedding_1"] = m.entity_embedding.detach().numpy().copy()
        out[name + "/relation_embedding_1"] = m.relation_embedding.detach().numpy().copy()
    np.savez_compressed(os.path.join(HERE, "oracle_models.npz"), **out)
    print("wrote oracle_models.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
