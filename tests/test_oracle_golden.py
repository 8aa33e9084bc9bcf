"""CPU: the oracle against the golden vectors produced by the reference's own TF sources
(``oracle/run_reference_tf.py``), plus oracle self-consistency (fp64 twin, cross-mode)."""
import os

import numpy as np
import pytest
import torch

from oracle import kge_oracle as ko

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# The reference's own acceptance tolerance for a train step (Checker/trainer_checker.py:91)
RTOL = ATOL = 1e-5


def _load(name):
    return np.load(os.path.join(GOLDEN, name))


def _interht_from_golden(g, ent_key="entity_embedding", rel_key="relation_embedding"):
    m = ko.OracleKGEModel("InterHT", int(g["E"]), int(g["R"]), int(g["H"]), float(g["gamma"]),
                          double_entity_embedding=True, triple_relation_embedding=True)
    with torch.no_grad():
        m.entity_embedding.copy_(torch.from_numpy(g[ent_key]))
        m.relation_embedding.copy_(torch.from_numpy(g[rel_key]))
    return m


def test_interht_scores_match_reference_source():
    g = _load("ref_tf_interht.npz")
    m = _interht_from_golden(g)
    pos = torch.from_numpy(g["positive_sample"])
    neg = torch.from_numpy(g["negative_sample"])
    with torch.no_grad():
        np.testing.assert_allclose(m(pos, "single").numpy(), g["score_single"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(m((pos, neg), "head-batch").numpy(), g["score_head_batch"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(m((pos, neg), "tail-batch").numpy(), g["score_tail_batch"], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("mode", [0, 1, 3])
def test_tf_call_matches_reference_source(mode):
    g = _load("ref_tf_interht.npz")
    m = _interht_from_golden(g)
    pos = torch.from_numpy(g["positive_sample"])
    neg = torch.from_numpy(g["negative_sample"])
    with torch.no_grad():
        got = ko.tf_call(m, (pos, neg), mode).numpy()
    np.testing.assert_allclose(got, g["call_mode%d" % mode], rtol=RTOL, atol=ATOL)


def test_tf_train_step_matches_reference_source():
    g = _load("ref_tf_trainstep.npz")
    m = _interht_from_golden(g, "entity_embedding_0", "relation_embedding_0")
    opt = ko.KerasAdam([m.entity_embedding, m.relation_embedding], lr=float(g["lr"]))
    for step in range(3):
        batch = tuple(torch.from_numpy(g["%s_%d" % (k, step)]) for k in
                      ("positive_sample", "negative_sample", "subsampling_weight", "mode"))
        loss = ko.tf_train_step(m, opt, batch)
        np.testing.assert_allclose(loss.item(), float(g["loss_%d" % step]), rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(m.entity_embedding.detach().numpy(), g["entity_embedding_%d" % (step + 1)],
                                   rtol=1e-4, atol=1e-6)
        np.testing.assert_allclose(m.relation_embedding.detach().numpy(), g["relation_embedding_%d" % (step + 1)],
                                   rtol=1e-4, atol=1e-6)


def test_loss_matches_reference_source():
    g = _load("ref_tf_loss.npz")
    neg = torch.from_numpy(g["negative_score"])
    pos = torch.from_numpy(g["positive_score"])
    w = torch.from_numpy(g["subsampling_weight"])

    class _M:
        entity_embedding = torch.from_numpy(g["entity_embedding"])
        relation_embedding = torch.from_numpy(g["relation_embedding"])

    for adv, uni, reg, temp, want in g["cases"]:
        _, _, loss = ko.kge_loss(neg, pos, w, negative_adversarial_sampling=bool(adv),
                                 adversarial_temperature=float(temp), uni_weight=bool(uni))
        if reg != 0.0:
            loss = loss + ko.l3_regularization(_M, float(reg))
        np.testing.assert_allclose(loss.item(), want, rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
def test_oracle_drift_detector(name):
    """Unpinned rows: oracle == what it produced when written (tests/golden/make_oracle_golden.py)."""
    g = _load("oracle_models.npz")
    de, dr, tr = ko.model_flags(name)
    ent = g[name + "/entity_embedding"]
    rel = g[name + "/relation_embedding"]
    H = ent.shape[1] // (2 if de else 1)
    gamma = {"TransE": 12.0, "DistMult": 12.0, "ComplEx": 12.0, "RotatE": 9.0, "pRotatE": 9.0, "InterHT": 24.0}[name]
    m = ko.OracleKGEModel(name, ent.shape[0], rel.shape[0], H, gamma, de, dr, tr)
    with torch.no_grad():
        m.entity_embedding.copy_(torch.from_numpy(ent))
        m.relation_embedding.copy_(torch.from_numpy(rel))
    pos = torch.from_numpy(g[name + "/positive_sample"])
    neg = torch.from_numpy(g[name + "/negative_sample"])
    w = torch.from_numpy(g[name + "/subsampling_weight"])
    with torch.no_grad():
        for mode in ko.MODES:
            sample = pos if mode == "single" else (pos, neg)
            np.testing.assert_allclose(m(sample, mode).numpy(), g[name + "/" + mode], rtol=1e-5, atol=1e-5)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=1e-2)
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=0.5)
    log = ko.train_step(m, opt, iter([(pos, neg, w, "tail-batch")]), args)
    np.testing.assert_allclose(log["loss"], float(g[name + "/loss"]), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(m.entity_embedding.detach().numpy(), g[name + "/entity_embedding_1"], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
def test_fp32_vs_fp64_twin(name):
    m = ko.make_model(name, 50, 4, 32, 9.0, seed=3)
    d = ko.to_double(m)
    pos, neg, _, _ = ko.synthetic_batch(50, 4, 5, 9, seed=1)
    with torch.no_grad():
        for mode in ko.MODES:
            sample = pos if mode == "single" else (pos, neg)
            np.testing.assert_allclose(m(sample, mode).numpy(), d(sample, mode).numpy(), rtol=2e-5, atol=2e-5)


@pytest.mark.parametrize("name", ko.MODEL_NAMES)
def test_cross_mode_consistency(name):
    """A negative equal to the true entity scores like 'single' (same triple, either batch mode)."""
    m = ko.make_model(name, 40, 3, 16, 9.0, seed=8)
    pos, _, _, _ = ko.synthetic_batch(40, 3, 6, 4, seed=2)
    with torch.no_grad():
        single = m(pos, "single")
        hb = m((pos, pos[:, 0:1]), "head-batch")
        tb = m((pos, pos[:, 2:3]), "tail-batch")
    np.testing.assert_allclose(hb.numpy(), single.numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(tb.numpy(), single.numpy(), rtol=1e-5, atol=1e-5)


def test_filter_rows_fast_equals_literal():
    triples = [(0, 0, 1), (2, 0, 1), (3, 0, 1), (0, 0, 4), (0, 1, 1), (5, 1, 6)]
    hr2t, rt2h = ko.build_true_maps(triples)
    for t in triples:
        for mode in ("head-batch", "tail-batch"):
            a = ko.test_dataset_row(t, set(triples), 8, mode)
            b = ko.test_dataset_row_fast(t, hr2t, rt2h, 8, mode)
            assert torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])


def test_validation_errors():
    with pytest.raises(ValueError):
        ko.OracleKGEModel("RotatE", 4, 2, 4, 1.0)
    with pytest.raises(ValueError):
        ko.OracleKGEModel("ComplEx", 4, 2, 4, 1.0, double_entity_embedding=True)
    with pytest.raises(ValueError):
        ko.OracleKGEModel("Nope", 4, 2, 4, 1.0)
    m = ko.make_model("TransE", 4, 2, 4, 1.0)
    with pytest.raises(ValueError):
        m(torch.zeros(1, 3, dtype=torch.int64), "bad-mode")


# ---- TranSparse (tensorflow_codes/model.py:96-106,226-235) ---------------------------------------------
def _transparse_from_golden(g, suffix=""):
    m = ko.OracleTranSparse(int(g["E"]), int(g["R"]), int(g["H"]), float(g["gamma"]))
    with torch.no_grad():
        m.entity_embedding.copy_(torch.from_numpy(g["entity_embedding" + suffix]))
        m.relation_embedding.copy_(torch.from_numpy(g["relation_embedding" + suffix]))
        m.W.copy_(torch.from_numpy(g["W" + suffix]))
        m.mask.copy_(torch.from_numpy(g["mask"]))
    return m


def test_transparse_scores_and_call_match_reference_source():
    g = _load("ref_tf_transparse.npz")
    m = _transparse_from_golden(g)
    pos, neg = torch.from_numpy(g["positive_sample"]), torch.from_numpy(g["negative_sample"])
    with torch.no_grad():
        np.testing.assert_allclose(m(pos, "single").numpy(), g["score_single"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(m((pos, neg), "head-batch").numpy(), g["score_head_batch"], rtol=RTOL, atol=ATOL)
        tb = m((pos, neg), "tail-batch").numpy()
        assert tb.shape == g["score_tail_batch"].shape == (pos.shape[0], 1)  # model.py:229: the tail is never read
        np.testing.assert_allclose(tb, g["score_tail_batch"], rtol=RTOL, atol=ATOL)
        for mode in (0, 1, 3):
            np.testing.assert_allclose(ko.tf_call(m, (pos, neg), mode).numpy(), g["call_mode%d" % mode], rtol=RTOL, atol=ATOL)


def test_transparse_train_steps_match_reference_source():
    g = _load("ref_tf_transparse.npz")
    m = _transparse_from_golden(g)
    opt = ko.KerasAdam([m.entity_embedding, m.relation_embedding, m.W], lr=float(g["lr"]))
    for step in range(3):
        batch = tuple(torch.from_numpy(g["%s_%d" % (k, step)]) for k in
                      ("positive_sample", "negative_sample", "subsampling_weight", "mode"))
        loss = ko.tf_train_step(m, opt, batch)
        np.testing.assert_allclose(loss.item(), float(g["loss_%d" % step]), rtol=RTOL, atol=ATOL)
        for k in ("entity_embedding", "relation_embedding", "W"):
            np.testing.assert_allclose(getattr(m, k).detach().numpy(), g["%s_%d" % (k, step + 1)], rtol=RTOL, atol=ATOL)
    # masked-out weights never move (their gradient is mask * dM = 0 and Adam's update of a zero history is 0)
    off = g["mask"] == 0
    assert np.array_equal(m.W.detach().numpy()[off], g["W"][off])


def test_transparse_projected_tail_variant_reads_the_tail():
    """project_tail=True (what model.py:229 evidently meant): the tail-batch scores depend on the candidates."""
    torch.manual_seed(3)
    m = ko.OracleTranSparse(31, 3, 12, 9.0, project_tail=True)
    pos, neg, _, _ = ko.synthetic_batch(31, 3, 4, 6, seed=1)
    with torch.no_grad():
        s = m((pos, neg), "tail-batch")
    assert s.shape == (4, 6) and float(s.std(dim=1).min()) > 0
