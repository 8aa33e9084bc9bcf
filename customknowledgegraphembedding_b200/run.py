"""``python -m customknowledgegraphembedding_b200.run`` -- the training / evaluation entry point with the flags of
the reference's two ``run.py`` files, wired to the B200 hot path.

* PyTorch codes (absent submodule): ``--do_train --do_valid --do_test --evaluate_train
  --countries --regions --data_path --model -de -dr -n -d -g -adv -a -b -r --test_batch_size --uni_weight -lr -cpu
  -init -save --max_steps --warm_up_steps --save_checkpoint_steps --valid_steps --log_steps --test_log_steps
  --nentity --nrelation``
* TF codes (/root/reference/tensorflow_codes/run.py:19-36): ``-ip/--input_path`` (TFRecord shards, one record per
  batch), ``-sf/--score_function``, ``-tr``, ``-bz``, ``--nentity``, ``--nrelation``, ``--hidden_dim``, ``--gamma``,
  ``--epochs``, ``--steps_per_epoch``.  A command with ``-ip`` and none of ``--do_train/--do_valid/--do_test/-init`` is
  the TF program's command line (README.md:10-20) and runs ``run_tf``: the TF trainer's semantics (gradient through the
  softmax weights, Keras Adam eps 1e-7, ``lrfn`` per-epoch schedule, ``steps_per_tpu_call = 99`` accounting).

What differs from upstream is only where the work happens: batches come from the on-GPU sampler (or from TFRecord
shards through the TF-free reader), ``train_step`` / ``test_step`` are the fused kernels.  This is glue -- no
arithmetic lives here.
"""
from __future__ import annotations

import argparse
import json
import logging
import os

from . import checkpoint as ck
from .model import KGEModel
from .sampler import BidirectionalGpuIterator, TrainTripleIndex
from .wire import TFRecordBatchReader


def parse_args(args=None):
    p = argparse.ArgumentParser(description="Training and Testing Knowledge Graph Embedding Models (B200 hot path)",
                                usage="run.py [<args>] [-h | --help]")
    p.add_argument("--cuda", action="store_true", help="accepted for compatibility: the kernels always run on CUDA")
    p.add_argument("--do_train", action="store_true")
    p.add_argument("--do_valid", action="store_true")
    p.add_argument("--do_test", action="store_true")
    p.add_argument("--evaluate_train", action="store_true", help="Evaluate on training data")
    p.add_argument("--countries", action="store_true", help="Use Countries S1/S2/S3 datasets")
    p.add_argument("--regions", type=int, nargs="+", default=None, help="Region Id for Countries S1/S2/S3 datasets, DO NOT MANUALLY SET")
    p.add_argument("--data_path", type=str, default=None)
    p.add_argument("-ip", "--input_path", type=str, default=None, help="TFRecord shard(s) / glob: one record per batch")
    p.add_argument("--model", "-sf", "--score_function", dest="model", default="TransE", type=str)
    p.add_argument("-de", "--double_entity_embedding", action="store_true")
    p.add_argument("-dr", "--double_relation_embedding", action="store_true")
    p.add_argument("-tr", "--triple_relation_embedding", action="store_true")
    p.add_argument("-n", "--negative_sample_size", default=128, type=int)
    p.add_argument("-d", "--hidden_dim", default=500, type=int)
    p.add_argument("-g", "--gamma", default=12.0, type=float)
    p.add_argument("-adv", "--negative_adversarial_sampling", action="store_true")
    p.add_argument("-a", "--adversarial_temperature", default=1.0, type=float)
    p.add_argument("-b", "-bz", "--batch_size", dest="batch_size", default=1024, type=int)
    p.add_argument("-r", "--regularization", default=0.0, type=float)
    p.add_argument("--test_batch_size", default=4, type=int, help="valid/test batch size")
    p.add_argument("--uni_weight", action="store_true", help="Otherwise use subsampling weighting like in word2vec")
    p.add_argument("-lr", "--learning_rate", default=0.0001, type=float)
    p.add_argument("-cpu", "--cpu_num", default=10, type=int, help="accepted for compatibility: no CPU workers are used")
    p.add_argument("-init", "--init_checkpoint", default=None, type=str)
    p.add_argument("-save", "--save_path", default=None, type=str)
    p.add_argument("--max_steps", default=100000, type=int)
    p.add_argument("--warm_up_steps", default=None, type=int)
    p.add_argument("--save_checkpoint_steps", default=10000, type=int)
    p.add_argument("--valid_steps", default=10000, type=int)
    p.add_argument("--log_steps", default=100, type=int, help="train log every xx steps")
    p.add_argument("--test_log_steps", default=1000, type=int, help="valid/test log every xx steps")
    p.add_argument("--nentity", type=int, default=0, help="DO NOT MANUALLY SET (needed with --input_path)")
    p.add_argument("--nrelation", type=int, default=0, help="DO NOT MANUALLY SET (needed with --input_path)")
    p.add_argument("--seed", type=int, default=0)
    # the TF trainer's own flags (/root/reference/tensorflow_codes/run.py:29-30)
    p.add_argument("--epochs", type=int, default=1)
    p.add_argument("--steps_per_epoch", type=int, default=1000)
    return p.parse_args(args)


def is_tf_command(args) -> bool:
    """The reference's TF entry point has no --do_train / -save (tensorflow_codes/run.py:20-37): a command that names
    TFRecord input and none of the PyTorch-side mode flags is that program's command line (README.md:10-20)."""
    return bool(args.input_path) and not (args.do_train or args.do_valid or args.do_test or args.init_checkpoint)


def run_tf(args, strategy=None):
    """``run(strategy, args)`` of /root/reference/tensorflow_codes/run.py:87-128 on the B200 kernels: TFRecord batches,
    ``TFKGEModel``, Keras-Adam with ``LRSchedule(step) = lrfn(step // steps_per_epoch)``, a ``Sum`` metric, and
    ``Trainer.training(steps_per_tpu_call=99, epochs, steps_per_epoch)`` -- gradient through the softmax weights,
    epsilon 1e-7, the reference's step accounting.  Returns the trainer."""
    from .tf_compat import KerasAdam, SumMetric, TFKGEModel, Trainer, _DefaultStrategy
    strategy = strategy if strategy is not None else _DefaultStrategy()
    print("Number of accelerators: ", strategy.num_replicas_in_sync)
    dataset = TFRecordBatchReader(args.input_path, args.batch_size, repeat=True, pin=True,
                                  nentity=args.nentity, nrelation=args.nrelation)
    kge_model = TFKGEModel(model_name=args.model, nentity=args.nentity, nrelation=args.nrelation,
                           hidden_dim=args.hidden_dim, gamma=args.gamma,
                           double_entity_embedding=args.double_entity_embedding,
                           double_relation_embedding=args.double_relation_embedding,
                           triple_relation_embedding=args.triple_relation_embedding)
    spe = args.steps_per_epoch
    optimizer = KerasAdam(learning_rate=lambda step: ck.lrfn(step // spe, strategy.num_replicas_in_sync))
    trainer = Trainer(strategy=strategy, dataloader=dataset, model=kge_model, optimizer=optimizer,
                      metrics=SumMetric("training_loss"))
    trainer.training(steps_per_tpu_call=99, epochs=args.epochs, steps_per_epoch=spe)
    return trainer


def read_triple(file_path, entity2id, relation2id):
    """Read triples and map them into ids (same as /root/reference/compress_data/main.py:30-36)."""
    triples = []
    with open(file_path) as fin:
        for line in fin:
            h, r, t = line.strip().split("\t")
            triples.append((entity2id[h], relation2id[r], entity2id[t]))
    return triples


def read_dicts(data_path):
    """entities.dict / relations.dict (compress_data/main.py:40-50)."""
    out = []
    for name in ("entities.dict", "relations.dict"):
        d = {}
        with open(os.path.join(data_path, name)) as fin:
            for line in fin:
                i, x = line.strip().split("\t")
                d[x] = int(i)
        out.append(d)
    return out


def log_metrics(mode, step, metrics):
    for metric in metrics:
        logging.info("%s %s at step %d: %f" % (mode, metric, step, metrics[metric]))


def main(args=None):
    args = parse_args(args) if not isinstance(args, argparse.Namespace) else args
    if is_tf_command(args):
        if not args.nentity or not args.nrelation:
            raise ValueError("--nentity / --nrelation are required with --input_path")
        return run_tf(args).model.kge, {}
    if (not args.do_train) and (not args.do_valid) and (not args.do_test):
        raise ValueError("one of train/val/test mode must be choosed.")
    if args.init_checkpoint:
        ck.override_config(args)
    elif args.data_path is None and args.input_path is None:
        raise ValueError("one of init_checkpoint/data_path must be choosed.")
    if args.do_train and args.save_path is None:
        raise ValueError("Where do you want to save your trained model?")
    if args.save_path:
        os.makedirs(args.save_path, exist_ok=True)
    logging.basicConfig(format="%(asctime)s %(levelname)-8s %(message)s", level=logging.INFO)

    train_triples = valid_triples = test_triples = all_true_triples = None
    if args.data_path:
        entity2id, relation2id = read_dicts(args.data_path)
        if args.countries:  # regions for Countries S1/S2/S3 datasets
            with open(os.path.join(args.data_path, "regions.list")) as fin:
                args.regions = [entity2id[line.strip()] for line in fin]
        args.nentity, args.nrelation = len(entity2id), len(relation2id)
        rd = lambda f: read_triple(os.path.join(args.data_path, f), entity2id, relation2id) \
            if os.path.exists(os.path.join(args.data_path, f)) else []
        train_triples, valid_triples, test_triples = rd("train.txt"), rd("valid.txt"), rd("test.txt")
        logging.info("#entity: %d  #relation: %d  #train: %d  #valid: %d  #test: %d" % (
            args.nentity, args.nrelation, len(train_triples), len(valid_triples), len(test_triples)))
        all_true_triples = train_triples + valid_triples + test_triples
    if not args.nentity or not args.nrelation:
        raise ValueError("--nentity / --nrelation are required without --data_path")

    kge_model = KGEModel(model_name=args.model, nentity=args.nentity, nrelation=args.nrelation,
                         hidden_dim=args.hidden_dim, gamma=args.gamma,
                         double_entity_embedding=args.double_entity_embedding,
                         double_relation_embedding=args.double_relation_embedding,
                         triple_relation_embedding=args.triple_relation_embedding).cuda()
    for name, param in kge_model.named_parameters():
        logging.info("Parameter %s: %s, require_grad = %s" % (name, str(param.size()), str(param.requires_grad)))

    init_step, sched = 0, None
    if args.do_train:
        if args.input_path:  # the TF trainer's input: pre-built batches (run.py:87-92)
            train_iterator = iter(TFRecordBatchReader(args.input_path, args.batch_size, repeat=True, pin=True,
                                                      as_torch_batch=True, nentity=args.nentity,
                                                      nrelation=args.nrelation))
        else:
            index = TrainTripleIndex(train_triples, args.nentity, args.nrelation)
            train_iterator = BidirectionalGpuIterator(index, args.negative_sample_size, args.batch_size, seed=args.seed)
        sched = ck.WarmupDecay(kge_model, args.learning_rate, args.max_steps, args.warm_up_steps)
    if args.init_checkpoint:
        logging.info("Loading checkpoint %s..." % args.init_checkpoint)
        saved = ck.load_checkpoint(kge_model, sched.optimizer if sched else None, args.init_checkpoint,
                                   do_train=args.do_train, map_location="cuda")
        init_step = saved.get("step", 0)
        if args.do_train:
            sched.current_learning_rate = saved["current_learning_rate"]
            sched.warm_up_steps = saved["warm_up_steps"]
    step = init_step

    if args.do_train:
        logging.info("learning_rate = %g" % sched.current_learning_rate)
        training_logs = []
        optimizer = sched.optimizer
        for step in range(init_step, args.max_steps):
            training_logs.append(kge_model.train_step(kge_model, optimizer, train_iterator, args))
            if step >= sched.warm_up_steps:
                optimizer = sched.after_step(step)
                logging.info("Change learning_rate to %g at step %d" % (sched.current_learning_rate, step))
            save_vars = {"step": step, "current_learning_rate": sched.current_learning_rate,
                         "warm_up_steps": sched.warm_up_steps}
            if step % args.save_checkpoint_steps == 0:
                ck.save_model(kge_model, optimizer, save_vars, args)
            if step % args.log_steps == 0:
                metrics = {m: sum(log[m] for log in training_logs) / len(training_logs) for m in training_logs[0]}
                log_metrics("Training average", step, metrics)
                training_logs = []
            if args.do_valid and valid_triples and step % args.valid_steps == 0:
                log_metrics("Valid", step, kge_model.test_step(kge_model, valid_triples, all_true_triples, args))
        ck.save_model(kge_model, optimizer, {"step": step, "current_learning_rate": sched.current_learning_rate,
                                             "warm_up_steps": sched.warm_up_steps}, args)

    results = {}
    for flag, name, triples in ((args.do_valid, "Valid", valid_triples), (args.do_test, "Test", test_triples),
                                (args.evaluate_train, "Train", train_triples)):
        if flag and triples:
            results[name] = kge_model.test_step(kge_model, triples, all_true_triples, args)
            log_metrics(name, step, results[name])
    if args.save_path and results:
        with open(os.path.join(args.save_path, "metrics.json"), "w") as f:
            json.dump(results, f)
    return kge_model, results


if __name__ == "__main__":
    main()
