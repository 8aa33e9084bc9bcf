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
