"""run.py: the reference's command-line surface (both flag sets) and one end-to-end train -> save -> reload -> test."""
import json
import os

import numpy as np
import pytest
import torch

from customknowledgegraphembedding_b200 import run


def test_flags_of_the_pytorch_codes():
    a = run.parse_args("--do_train --cuda --do_valid --do_test --data_path data/FB15k-237 --model RotatE -n 256 -b 1024 "
                       "-d 1000 -g 24.0 -a 1.0 -adv -lr 0.0001 --max_steps 150000 -save models/x --test_batch_size 16 "
                       "-de".split())
    assert (a.model, a.negative_sample_size, a.batch_size, a.hidden_dim, a.gamma) == ("RotatE", 256, 1024, 1000, 24.0)
    assert a.negative_adversarial_sampling and a.double_entity_embedding and not a.double_relation_embedding
    assert (a.adversarial_temperature, a.learning_rate, a.max_steps, a.test_batch_size) == (1.0, 1e-4, 150000, 16)
    assert a.warm_up_steps is None and a.regularization == 0.0 and not a.uni_weight


def test_flags_of_the_tf_codes():
    # /root/reference/tensorflow_codes/run.py:19-36
    a = run.parse_args("-ip shards/*.tfrec -sf InterHT -tr -de -bz 16 --nentity 40943 --nrelation 11 --hidden_dim 1000 "
                       "--gamma 24.0 --do_train -save out".split())
    assert (a.input_path, a.model, a.batch_size, a.nentity, a.nrelation) == ("shards/*.tfrec", "InterHT", 16, 40943, 11)
    assert a.triple_relation_embedding and a.double_entity_embedding


README_COMMAND = ("-ip gs://hien7613storage2/datasets/KGE/wn18rr.tfrec -bz 16 -sf InterHT --nentity 40943 --nrelation 11 "
                  "--hidden_dim 1000 --gamma 24.0 --epochs 10 --steps_per_epoch 100 -de -tr")


def test_readme_command_of_the_reference_verbatim():
    """/root/reference/README.md:10-20 -- the only documented command line: it must parse as is and route to the TF
    trainer (no --do_train / -save exist in tensorflow_codes/run.py:20-37)."""
    a = run.parse_args(README_COMMAND.split())
    assert run.is_tf_command(a)
    assert (a.model, a.batch_size, a.nentity, a.nrelation, a.hidden_dim, a.gamma) == ("InterHT", 16, 40943, 11, 1000, 24.0)
    assert (a.epochs, a.steps_per_epoch) == (10, 100) and a.double_entity_embedding and a.triple_relation_embedding
    # defaults of tensorflow_codes/run.py:29-30
    b = run.parse_args("-ip x.tfrec -bz 4 -sf InterHT --nentity 9 --nrelation 2 --hidden_dim 4 --gamma 1 -de -tr".split())
    assert (b.epochs, b.steps_per_epoch) == (1, 1000)
    # a PyTorch-side mode flag keeps the upstream loop
    assert not run.is_tf_command(run.parse_args("-ip x.tfrec --do_train -save out --nentity 9 --nrelation 2".split()))


@pytest.mark.gpu
def test_readme_command_runs_the_tf_trainer(tmp_path, capsys):
    """The README command shape on a small shard: Trainer.training's accounting (one real step counted as 99;
    supervisor.py:40-42), Keras Adam on the lrfn schedule, gradient through the softmax -- first-step loss against the
    oracle's restatement of supervisor.py:15-28."""
    from oracle import kge_oracle as ko
    from customknowledgegraphembedding_b200 import wire
    E, R, H, B, N = 120, 3, 16, 8, 12
    batches = []
    for s in range(4):
        pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=s)
        batches.append((pos.numpy(), neg.numpy(), w.numpy().reshape(-1, 1), 0 if mode == "head-batch" else 1))
    shard = str(tmp_path / "toy.tfrec")
    assert wire.write_tfrecords(batches, shard, B) == 4
    torch.manual_seed(0)
    cmd = "-ip %s -bz %d -sf InterHT --nentity %d --nrelation %d --hidden_dim %d --gamma 24.0 --epochs 2 " \
          "--steps_per_epoch 100 -de -tr" % (shard, B, E, R, H)
    torch.manual_seed(0)
    om = ko.make_model("InterHT", E, R, H, 24.0, seed=0)
    import customknowledgegraphembedding_b200.tf_compat as tfc
    seen = {}
    orig = tfc.Trainer.train_step

    def spy(self, data_iter):
        if "w0" not in seen:
            seen["w0"] = {k: v.detach().cpu().clone() for k, v in self.model.state_dict().items()}
        out = orig(self, data_iter)
        seen.setdefault("losses", []).append(float(out["loss"]))
        seen["opt"] = self.optimizer
        return out
    tfc.Trainer.train_step = spy
    try:
        model, _ = run.main(cmd.split())
    finally:
        tfc.Trainer.train_step = orig
    # step += 99 per call, epoch = step // 100, stop at epoch >= 2: three real steps
    assert len(seen["losses"]) == 3 and seen["opt"].iterations == 3
    text = capsys.readouterr().out
    assert "EPOCH 1/2" in text and "DONE" in text
    om.load_state_dict(seen["w0"])
    opt = ko.KerasAdam([om.entity_embedding, om.relation_embedding], lr=1e-5)  # lrfn(0) = LR_START
    pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=0)
    want = ko.tf_train_step(om, opt, (pos, neg, w.view(-1, 1), torch.full((B,), 0 if mode == "head-batch" else 1)))
    np.testing.assert_allclose(seen["losses"][0], float(want), rtol=1e-5, atol=1e-6)
    with pytest.raises(IndexError):  # ids beyond --nentity: refused like tf.gather would
        run.main(cmd.replace("--nentity %d" % E, "--nentity 50").split())


@pytest.mark.gpu
def test_tf_command_with_the_transparse_score_function(tmp_path, capsys):
    """`-sf TranSparse` (tensorflow_codes/model.py:96-106,226-235) through the same command line: W and the mask are
    built, the trainer steps all three variables, the first loss equals the oracle's on the same variables."""
    from oracle import kge_oracle as ko
    from customknowledgegraphembedding_b200 import wire
    import customknowledgegraphembedding_b200.tf_compat as tfc
    E, R, H, B, N = 90, 3, 24, 8, 10
    batches = []
    for s in range(3):
        pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=s)
        batches.append((pos.numpy(), neg.numpy(), w.numpy().reshape(-1, 1), 0 if mode == "head-batch" else 1))
    shard = str(tmp_path / "toy.tfrec")
    assert wire.write_tfrecords(batches, shard, B) == 3
    cmd = "-ip %s -bz %d -sf TranSparse --nentity %d --nrelation %d --hidden_dim %d --gamma 12.0 --epochs 1 " \
          "--steps_per_epoch 100" % (shard, B, E, R, H)
    seen = {}
    orig = tfc.Trainer.train_step

    def spy(self, data_iter):
        if "w0" not in seen:
            seen["w0"] = {k: v.detach().cpu().clone() for k, v in self.model.state_dict().items()}
        out = orig(self, data_iter)
        seen.setdefault("losses", []).append(float(out["loss"]))
        return out
    tfc.Trainer.train_step = spy
    try:
        model, _ = run.main(cmd.split())
    finally:
        tfc.Trainer.train_step = orig
    assert len(seen["losses"]) == 2 and "DONE" in capsys.readouterr().out   # step += 99 per call: two real steps
    assert model.W.shape == (R, H, H) and not torch.equal(model.W.detach().cpu(), seen["w0"]["W"])
    om = ko.OracleTranSparse(E, R, H, 12.0)
    om.load_state_dict(seen["w0"])
    opt = ko.KerasAdam([om.entity_embedding, om.relation_embedding, om.W], lr=1e-5)
    pos, neg, w, mode = ko.synthetic_batch(E, R, B, N, seed=0)
    want = ko.tf_train_step(om, opt, (pos, neg, w.view(-1, 1), torch.full((B,), 0 if mode == "head-batch" else 1)))
    np.testing.assert_allclose(seen["losses"][0], float(want), rtol=1e-5, atol=1e-6)
    with pytest.raises(ValueError):  # model.py:227 needs entity_dim == relation_dim
        run.main((cmd + " -de").split())


def test_argument_errors_match_upstream():
    with pytest.raises(ValueError, match="one of train/val/test mode must be choosed."):
        run.main([])
    with pytest.raises(ValueError, match="one of init_checkpoint/data_path must be choosed."):
        run.main(["--do_test"])
    with pytest.raises(ValueError, match="Where do you want to save your trained model?"):
        run.main(["--do_train", "--data_path", "x"])


def _write_dataset(root, E=60, R=3, seed=0):
    """A learnable toy graph: relation k maps entity i to (i + k + 1) mod E."""
    rng = np.random.default_rng(seed)
    triples = [(h, r, (h + r + 1) % E) for h in range(E) for r in range(R)]
    rng.shuffle(triples)
    os.makedirs(root, exist_ok=True)
    with open(os.path.join(root, "entities.dict"), "w") as f:
        f.writelines("%d\te%d\n" % (i, i) for i in range(E))
    with open(os.path.join(root, "relations.dict"), "w") as f:
        f.writelines("%d\tr%d\n" % (i, i) for i in range(R))
    n_test = 12
    for name, part in (("train", triples[2 * n_test:]), ("valid", triples[:n_test]), ("test", triples[n_test:2 * n_test])):
        with open(os.path.join(root, name + ".txt"), "w") as f:
            f.writelines("e%d\tr%d\te%d\n" % t for t in part)


@pytest.mark.gpu
def test_train_save_reload_test(tmp_path):
    data, save = str(tmp_path / "data"), str(tmp_path / "model")
    _write_dataset(data)
    common = ["--data_path", data, "--model", "RotatE", "-de", "-d", "32", "-g", "6.0", "-n", "16", "-b", "32", "-adv",
              "-lr", "0.02", "--test_batch_size", "8", "-save", save]
    torch.manual_seed(0)
    model, res = run.main(["--do_train", "--do_test", "--max_steps", "600", "--log_steps", "200",
                           "--save_checkpoint_steps", "1000", "--valid_steps", "1000"] + common)
    # the rotation structure is learnable: filtered ranks on held-out triples are far better than chance (MR 30)
    assert res["Test"]["MRR"] > 0.5 and res["Test"]["HITS@10"] > 0.8, res
    for f in ("config.json", "checkpoint", "entity_embedding.npy", "relation_embedding.npy", "metrics.json"):
        assert os.path.exists(os.path.join(save, f)), f
    # -init: configuration comes from config.json, weights from the checkpoint; same metrics as before
    model2, res2 = run.main(["--do_test", "-init", save, "--test_batch_size", "8"])
    assert torch.equal(model2.entity_embedding.cpu(), model.entity_embedding.cpu())
    assert res2["Test"] == res["Test"]
    assert json.load(open(os.path.join(save, "config.json")))["model"] == "RotatE"


@pytest.mark.gpu
def test_resume_keeps_the_staged_fast_path(tmp_path):
    """-init with --do_train: Adam's `step` counters come back on the host (torch.load(map_location="cuda") would leave
    them on the GPU: a device sync per step and no argument cache), so the resumed run takes the staged step."""
    data, save = str(tmp_path / "data"), str(tmp_path / "model")
    _write_dataset(data)
    common = ["--data_path", data, "--model", "RotatE", "-de", "-d", "16", "-g", "6.0", "-n", "8", "-b", "16", "-adv",
              "-lr", "0.01", "-save", save, "--warm_up_steps", "1000", "--log_steps", "1000", "--save_checkpoint_steps", "1000", "--valid_steps", "1000"]
    run.main(["--do_train", "--max_steps", "5"] + common)
    model, _ = run.main(["--do_train", "-init", save, "--max_steps", "9"] + common)
    ck = torch.load(os.path.join(save, "checkpoint"), weights_only=False)
    assert ck["step"] == 8
    st = ck["optimizer_state_dict"]["state"]
    # 5 steps, then steps 4..8 of the resumed run; the counters were saved from host tensors
    assert all(int(s["step"]) == 10 and s["step"].device.type == "cpu" for s in st.values())
