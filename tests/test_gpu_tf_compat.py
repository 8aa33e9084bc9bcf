"""GPU: the TF-shaped twin (TFKGEModel / Trainer / KerasAdam) against golden vectors produced by running the
reference's own tensorflow_codes/model.py and supervisor.py (oracle/run_reference_tf.py, tests/golden/ref_tf_*.npz)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _model(g, ent_key, rel_key):
    from customknowledgegraphembedding_b200.tf_compat import TFKGEModel
    m = TFKGEModel("InterHT", int(g["E"]), int(g["R"]), int(g["H"]), float(g["gamma"]),
                   double_entity_embedding=True, triple_relation_embedding=True)
    with torch.no_grad():
        m.entity_embedding.copy_(torch.from_numpy(g[ent_key]))
        m.relation_embedding.copy_(torch.from_numpy(g[rel_key]))
    return m


@pytest.mark.parametrize("mode", [0, 1, 3])
def test_call_matches_reference_source(mode):
    g = np.load(os.path.join(GOLDEN, "ref_tf_interht.npz"))
    m = _model(g, "entity_embedding", "relation_embedding")
    pos = torch.from_numpy(g["positive_sample"]).cuda()
    neg = torch.from_numpy(g["negative_sample"]).cuda()
    with torch.no_grad():
        got = m(((pos, neg), torch.tensor(mode))).cpu().numpy()
    assert got.shape == g["call_mode%d" % mode].shape
    np.testing.assert_allclose(got, g["call_mode%d" % mode], rtol=1e-5, atol=1e-5)  # Checker/trainer_checker.py:91


def test_call_gradient_flows_through_softmax():
    """model.py:168-171 has no stop-gradient: d/dscore must include the softmax term (autograd through K1/K3)."""
    from oracle import kge_oracle as ko
    g = np.load(os.path.join(GOLDEN, "ref_tf_interht.npz"))
    m = _model(g, "entity_embedding", "relation_embedding")
    om = ko.OracleKGEModel("InterHT", int(g["E"]), int(g["R"]), int(g["H"]), float(g["gamma"]), True, False, True)
    om.load_state_dict(m.state_dict())
    pos, neg = torch.from_numpy(g["positive_sample"]), torch.from_numpy(g["negative_sample"])
    ko.tf_call(om, (pos, neg), 1).sum().backward()
    m(((pos.cuda(), neg.cuda()), 1)).sum().backward()
    want = om.entity_embedding.grad.numpy()
    got = m.entity_embedding.grad.cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=2e-5, atol=2e-5 * np.abs(want).max())


def test_trainer_train_step_matches_reference_source():
    from customknowledgegraphembedding_b200.tf_compat import KerasAdam, SumMetric, Trainer
    g = np.load(os.path.join(GOLDEN, "ref_tf_trainstep.npz"))
    m = _model(g, "entity_embedding_0", "relation_embedding_0")
    opt = KerasAdam(learning_rate=float(g["lr"]))
    metric = SumMetric()
    batches = [tuple(torch.from_numpy(g["%s_%d" % (k, s)]) for k in
                     ("positive_sample", "negative_sample", "subsampling_weight", "mode")) for s in range(3)]
    trainer = Trainer(None, batches, m, opt, metric)
    it = iter(batches)
    for step in range(3):
        metric.reset_states()
        out = trainer.train_step(it)
        np.testing.assert_allclose(float(out["loss"]), float(g["loss_%d" % step]), rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(metric.result(), float(g["loss_%d" % step]), rtol=1e-5, atol=1e-5)
        for name in ("entity_embedding", "relation_embedding"):
            want = g["%s_%d" % (name, step + 1)]
            got = getattr(m, name).detach().cpu().numpy()
            # Adam steps of size lr = 1e-2 on values ~2 (ulp 2.4e-7); an element whose gradient is within a few
            # roundings of zero moves by a visibly different fraction of lr in any two fp32 implementations
            np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-3 * float(g["lr"]))
            assert (np.abs(got - want) > 1e-4 * float(g["lr"])).mean() < 0.01


def test_zero_learning_rate_like_the_checker():
    """Checker/trainer_checker.py:10-14 sets both learning rates to 0: weights must not move."""
    from customknowledgegraphembedding_b200.tf_compat import KerasAdam, Trainer
    g = np.load(os.path.join(GOLDEN, "ref_tf_trainstep.npz"))
    m = _model(g, "entity_embedding_0", "relation_embedding_0")
    opt = KerasAdam(learning_rate=0.01)
    opt.learning_rate.assign(0.0)
    before = m.entity_embedding.detach().clone()
    batch = tuple(torch.from_numpy(g["%s_0" % k]) for k in ("positive_sample", "negative_sample", "subsampling_weight", "mode"))
    out = Trainer(None, None, m, opt, None).train_step(batch)
    assert torch.equal(before, m.entity_embedding.detach())
    np.testing.assert_allclose(float(out["loss"]), float(g["loss_0"]), rtol=1e-5, atol=1e-5)
