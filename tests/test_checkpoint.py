"""N3: checkpoint layout + learning-rate schedules of the reference (CPU: tensors only, no kernels)."""
import argparse
import json
import os

import numpy as np
import torch


def _model(name="RotatE"):
    import customknowledgegraphembedding_b200 as kb
    torch.manual_seed(0)
    return kb.KGEModel(name, 30, 4, 8, 9.0, True, False)


def test_save_load_layout_and_round_trip(tmp_path):
    from customknowledgegraphembedding_b200 import checkpoint as ck
    m = _model()
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=1e-3)
    for p in (m.entity_embedding, m.relation_embedding):   # the state layout the fused step keeps up to date
        st = m._adam_state(opt, p)
        st["step"] += 3
        st["exp_avg"].normal_()
        st["exp_avg_sq"].uniform_()
    args = argparse.Namespace(save_path=str(tmp_path / "ckpt"), model="RotatE", hidden_dim=8, gamma=9.0,
                              double_entity_embedding=True, double_relation_embedding=False, data_path="data/x",
                              test_batch_size=4, countries=False)
    ck.save_model(m, opt, {"step": 17, "current_learning_rate": 5e-5, "warm_up_steps": 50}, args)
    assert sorted(os.listdir(args.save_path)) == ["checkpoint", "config.json", "entity_embedding.npy", "relation_embedding.npy"]
    assert json.load(open(os.path.join(args.save_path, "config.json")))["hidden_dim"] == 8
    assert np.array_equal(np.load(os.path.join(args.save_path, "entity_embedding.npy")), m.entity_embedding.detach().numpy())
    raw = torch.load(os.path.join(args.save_path, "checkpoint"), weights_only=False)
    assert {"step", "current_learning_rate", "warm_up_steps", "model_state_dict", "optimizer_state_dict"} <= set(raw)
    assert {"entity_embedding", "relation_embedding", "gamma", "embedding_range"} <= set(raw["model_state_dict"])  # copy_weights.py:9-10

    torch.manual_seed(1)
    m2 = _model()
    with torch.no_grad():
        m2.entity_embedding.add_(1.0)
    opt2 = torch.optim.Adam(filter(lambda p: p.requires_grad, m2.parameters()), lr=1.0)
    saved = ck.load_checkpoint(m2, opt2, args.save_path)
    assert saved == {"step": 17, "current_learning_rate": 5e-5, "warm_up_steps": 50}
    assert torch.equal(m2.entity_embedding, m.entity_embedding) and torch.equal(m2.relation_embedding, m.relation_embedding)
    assert opt2.param_groups[0]["lr"] == 1e-3
    for p, p2 in ((m.entity_embedding, m2.entity_embedding), (m.relation_embedding, m2.relation_embedding)):
        assert torch.equal(opt.state[p]["exp_avg"], opt2.state[p2]["exp_avg"])
        assert float(opt2.state[p2]["step"]) == 3.0
    assert m2._fused_adam_plan(opt2) is not None   # the restored optimizer still qualifies for the fused step

    new = argparse.Namespace(init_checkpoint=args.save_path, data_path=None, model=None, hidden_dim=None,
                             double_entity_embedding=None, double_relation_embedding=None, test_batch_size=None)
    ck.override_config(new)
    assert (new.model, new.hidden_dim, new.double_entity_embedding, new.data_path) == ("RotatE", 8, True, "data/x")


def test_warmup_decay_matches_upstream_loop():
    from customknowledgegraphembedding_b200 import checkpoint as ck
    m = _model("TransE") if False else _model()
    sched = ck.WarmupDecay(m, learning_rate=1e-3, max_steps=100)
    assert sched.warm_up_steps == 50
    first = sched.optimizer
    lrs = []
    for step in range(160):
        opt = sched.after_step(step)
        lrs.append(opt.param_groups[0]["lr"])
    assert lrs[49] == 1e-3 and abs(lrs[50] - 1e-4) < 1e-18 and abs(lrs[150] - 1e-5) < 1e-18
    assert sched.warm_up_steps == 450 and sched.optimizer is not first and len(sched.optimizer.state) == 0


def test_lrfn_is_the_tf_schedule():
    from customknowledgegraphembedding_b200 import checkpoint as ck
    assert ck.lrfn(0) == 0.00001
    assert abs(ck.lrfn(2.5) - 0.00003) < 1e-12
    assert abs(ck.lrfn(5) - 0.00005) < 1e-12            # ramp-up over, no sustain: decay starts at LR_MAX
    assert abs(ck.lrfn(6) - (0.00004 * 0.8 + 0.00001)) < 1e-12
    assert abs(ck.lrfn(3, num_replicas_in_sync=8) - ((0.0004 - 0.00001) / 5 * 3 + 0.00001)) < 1e-12
