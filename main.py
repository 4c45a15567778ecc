#!/usr/bin/env python
"""Training entry point — same name, flags and roles as the reference's main.py, re-built for B200.

    python main.py                                                        # local mode (no --worker_hosts), 1 GPU or CPU
    python main.py --job_name=ps     --task_id=0 --ps_hosts=H:P --worker_hosts=H:P1,H:P2
    python main.py --job_name=worker --task_id=k --ps_hosts=H:P --worker_hosts=H:P1,H:P2 --train_dir=/shared/dir
    python main.py --num_gpus=8                                           # tower mode: one process per GPU on this node

Reference surface kept (main.py:10-32): --job_name, --ps_hosts, --worker_hosts, --task_id (alias --task_index),
--train_dir, --log_device_placement.  Hyper-parameters come from the parameter manager (the reference's
tf_parameter_mgr): a JSON file (--params_file / $DTM_PARAMS) that individual flags below may override.

What changed underneath: the reference's asynchronous gRPC parameter-server updates are replaced by
synchronous data parallelism — gradients are all-reduced inside a peer-memory kernel fused with the optimizer — so a
`ps` process only hosts the rendezvous store and blocks like ``server.join()`` until the workers are done.
"""
from __future__ import annotations

import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from distributed_tensorflow_ibm_mnist_b200.config import flags  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.config import parameter_mgr as tf_parameter_mgr  # noqa: E402

FLAGS = flags.FLAGS

flags.DEFINE_string("job_name", "", 'One of "ps", "worker"')
flags.DEFINE_string("ps_hosts", "", "Comma-separated list of hostname:port for the parameter server jobs")
flags.DEFINE_string("worker_hosts", "", "Comma-separated list of hostname:port for the worker jobs")
flags.DEFINE_integer("task_id", 0, "Task ID of the worker/replica running the training")
flags.DEFINE_alias("task_index", "task_id")
flags.DEFINE_string("train_dir", "/tmp/mnist_train", "Directory where to write event logs and checkpoint")
flags.DEFINE_boolean("log_device_placement", False, "Whether to log device placement (rank -> GPU, variable -> flat-buffer offsets)")
# ---- additions (the reference fixes these inside the absent tf_parameter_mgr module) --------------------------------
flags.DEFINE_string("params_file", "", "JSON file with hyper-parameters and data paths (see config/parameter_mgr.py)")
flags.DEFINE_string("model", "", "Model preset: reference | baseline | wide")
flags.DEFINE_integer("channels", 0, "Input channels (1 or 3); 0 = preset default")
flags.DEFINE_integer("batch_size", 0, "Per-worker batch size (0 = parameter manager)")
flags.DEFINE_integer("max_steps", 0, "Global step budget (0 = parameter manager)")
flags.DEFINE_integer("test_interval", 0, "Steps between summary dumps (0 = parameter manager)")
flags.DEFINE_float("base_lr", 0.0, "Base learning rate (0 = parameter manager)")
flags.DEFINE_float("lr_decay", 0.0, "Learning-rate decay factor (0 = parameter manager)")
flags.DEFINE_string("optimizer", "", "sgd | momentum | adam ('' = parameter manager)")
flags.DEFINE_string("train_data", "", "Comma-separated TFRecord files for training")
flags.DEFINE_string("test_data", "", "Comma-separated TFRecord files for the test split")
flags.DEFINE_integer("input_threads", 2, "Reader threads of the TFRecord input pipeline (mnist_input.py:68 fixes num_threads=2)")
flags.DEFINE_boolean("use_test_split", False, "Train on the test split, as the reference does (main.py:85)")
flags.DEFINE_string("engine", "auto", "auto | native (sm_100a kernels) | torch (portable oracle path)")
flags.DEFINE_string("backend", "", "torch.distributed backend for the control plane ('' = nccl on GPU, gloo on CPU)")
flags.DEFINE_integer("num_gpus", 0, "Tower mode: k GPUs of this node driven by ONE process, one thread and one engine per GPU "
                                    "(0 = off); see --tower_procs")
flags.DEFINE_boolean("tower_procs", False, "Tower mode with one worker PROCESS per GPU (the only form for --engine=torch)")
flags.DEFINE_integer("max_restarts", 0, "Supervised run: restart the worker set from the newest checkpoint up to this many times "
                                        "when a worker fails (the recoverable session of MonitoredTrainingSession)")
flags.DEFINE_boolean("verbose_steps", False, 'Print "training" and the step every iteration like the reference (main.py:149-150)')
flags.DEFINE_float("save_checkpoint_secs", -1.0, "Seconds between checkpoints (-1 = parameter manager, default 600)")
flags.DEFINE_integer("seed", -1, "Random seed (-1 = parameter manager)")
flags.DEFINE_boolean("no_cuda_graph", False, "Launch kernels individually instead of replaying a CUDA graph")


def _configure_hyper_params():
    if FLAGS.params_file:
        tf_parameter_mgr.load_file(FLAGS.params_file)
    over = {}
    if FLAGS.model:
        over["model"] = FLAGS.model
    if FLAGS.channels:
        over["channels"] = FLAGS.channels
    if FLAGS.batch_size:
        over["train_batch_size"] = FLAGS.batch_size
    if FLAGS.max_steps:
        over["max_steps"] = FLAGS.max_steps
    if FLAGS.test_interval:
        over["test_interval"] = FLAGS.test_interval
    if FLAGS.base_lr:
        over["base_learning_rate"] = FLAGS.base_lr
    if FLAGS.lr_decay:
        over["learning_rate_decay"] = FLAGS.lr_decay
    if FLAGS.optimizer:
        over["optimizer"] = FLAGS.optimizer
    if FLAGS.train_data:
        over["train_data"] = FLAGS.train_data
    if FLAGS.test_data:
        over["test_data"] = FLAGS.test_data
    if FLAGS.save_checkpoint_secs >= 0:
        over["save_checkpoint_secs"] = FLAGS.save_checkpoint_secs
    if FLAGS.seed >= 0:
        over["seed"] = FLAGS.seed
    return tf_parameter_mgr.configure(**over)


def setup_distribute():
    """Reference main.py:42-68: build the cluster description and this process's role; ps processes never return."""
    from distributed_tensorflow_ibm_mnist_b200.cluster import spec as cluster_spec

    cluster = cluster_spec.ClusterSpec.from_flags(FLAGS.ps_hosts, FLAGS.worker_hosts)
    if cluster.is_local:
        FLAGS.task_id = 0
        return cluster, cluster_spec.resolve_role(cluster, "", 0)
    print("Cluster spec: ", cluster.as_dict())
    role = cluster_spec.resolve_role(cluster, FLAGS.job_name, FLAGS.task_id)
    if role.job_name == "ps":
        cluster_spec.serve_parameter_server(cluster, role)   # like server.join(), but returns when the workers finish
        return cluster, None
    return cluster, role


def train(role=None, plane=None, hp=None):
    """One worker.  ``role``/``plane`` are given by the single-process tower mode (the worker is a thread, its control plane a
    ThreadPlane); otherwise the role comes from the flags and the control plane is a torch.distributed process group."""
    import torch

    from distributed_tensorflow_ibm_mnist_b200.cluster import spec as cluster_spec
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import DeviceFeeder, SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.formats.events import EventFileWriter
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing, monitor as monitor_cb, trainer

    threaded = plane is not None
    if hp is None:
        hp = _configure_hyper_params()
    if threaded:
        cluster = cluster_spec.ClusterSpec()
    else:
        cluster, role = setup_distribute()
        if role is None:
            return 0  # parameter-server process: done
    is_chief = role.is_chief
    max_steps, test_interval, batch_size = (tf_parameter_mgr.getMaxSteps(), tf_parameter_mgr.getTestInterval(),
                                            tf_parameter_mgr.getTrainBatchSize())

    use_cuda = torch.cuda.is_available() and FLAGS.engine != "torch"
    if FLAGS.engine == "native" and not torch.cuda.is_available():
        raise SystemExit("--engine=native needs a CUDA device")
    local_rank = role.rank if threaded else int(os.environ.get("LOCAL_RANK", role.rank % max(1, torch.cuda.device_count()) if use_cuda else 0))
    device = torch.device("cuda", local_rank) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)

    group = store = comm = None
    if threaded:
        from distributed_tensorflow_ibm_mnist_b200.parallel.peer_comm import PeerComm

        group = plane
        comm = PeerComm(plane, device)
    elif role.world_size > 1:
        backend = FLAGS.backend or ("nccl" if use_cuda else "gloo")
        group, store = cluster_spec.init_worker_group(cluster, role, backend)
        if use_cuda:
            from distributed_tensorflow_ibm_mnist_b200.parallel.peer_comm import PeerComm

            comm = PeerComm(group, device)

    spec = presets.make_spec(hp.model, hp.channels)
    opt = tf_parameter_mgr.getOptimizer(tf_parameter_mgr.getBaseLearningRate())
    kw = dict(optimizer=opt.kind, base_lr=tf_parameter_mgr.getBaseLearningRate(), decay_rate=tf_parameter_mgr.getLearningRateDecay(),
              decay_steps=presets.decay_steps(batch_size), momentum=opt.momentum, seed=hp.seed)
    if use_cuda:
        from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

        engine = NativeEngine(spec, batch_size, device=device, comm=comm, use_graph=not FLAGS.no_cuda_graph, **kw)
    else:
        from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

        engine = TorchEngine(spec, batch_size, device=device, group=group, **kw)

    if FLAGS.log_device_placement:
        print(f"[placement] worker {role.task_id} -> {device} ({type(engine).__name__}); world={role.world_size}")
        if hasattr(engine, "flat"):
            for v in spec.variables:
                print(f"[placement]   {v.name:28s} {str(v.shape):18s} flat offset {engine.flat.offsets[v.name]}")

    # restore-or-init (what MonitoredTrainingSession does for the chief; every worker reads the same files, the
    # subsequent broadcast makes rank 0's copy authoritative)
    os.makedirs(FLAGS.train_dir, exist_ok=True)
    step = checkpointing.restore_checkpoint(FLAGS.train_dir, engine, spec, opt.kind) if checkpointing.resolve_checkpoint(FLAGS.train_dir) else None
    if step is None:
        engine.load_params(oracle.init_params(spec, seed=hp.seed))
    else:
        print(f"restored checkpoint at global step {step} from {FLAGS.train_dir}")
    if comm is not None:
        engine.broadcast_params()
    elif group is not None:
        import torch.distributed as dist

        for t in engine.w.values():
            dist.broadcast(t, src=0, group=group)

    # input pipeline (mnist_input.inputs): real TFRecords when configured, otherwise synthetic MNIST-shaped batches
    files = tf_parameter_mgr.getTestData() if FLAGS.use_test_split else tf_parameter_mgr.getTrainData()
    if files:
        from distributed_tensorflow_ibm_mnist_b200.data.readers import TFRecordBatches

        source = TFRecordBatches(files, batch_size, spec.image_pixels, seed=hp.seed + role.rank, pin=use_cuda,
                                 num_threads=max(1, FLAGS.input_threads))
        print("image Shape")
        print((spec.image_pixels,))
    else:
        source = SyntheticMnist(batch_size, spec.image_pixels, pool_batches=16, seed=hp.seed + role.rank, separable=True,
                                pin=use_cuda)
    feeder = DeviceFeeder(engine, source) if use_cuda else None

    hooks = [trainer.StopAtStepHook(last_step=max_steps), trainer.NanLossHook()]
    if os.environ.get("DTM_FAULT_INJECT"):
        hooks.append(trainer.FaultInjectionHook(os.environ["DTM_FAULT_INJECT"], role.rank, int(os.environ.get("DTM_RESTART_COUNT", "0"))))
    saver = None
    if getattr(engine, "sharded", False):   # every rank helps assembling checkpoints; only the chief writes them
        saver = checkpointing.CheckpointSaver(FLAGS.train_dir, spec, opt.kind, save_secs=hp.save_checkpoint_secs,
                                              num_shards=max(1, cluster.num_ps), max_to_keep=hp.max_to_keep)
    if is_chief:
        # what MonitoredTrainingSession does on the chief before the first step: graph.pbtxt + a step-rate counter
        checkpointing.write_graph(FLAGS.train_dir, spec, opt.kind)
        hooks.append(trainer.StepCounterHook(EventFileWriter(FLAGS.train_dir), 100, batch_size * max(1, role.world_size)))
        log_dir = os.path.join(FLAGS.train_dir, "log")
        monitor = monitor_cb.CMonitor(log_dir, test_interval, max_steps)
        writer = EventFileWriter(log_dir)
        for layer in [l for l in ("conv1", "conv2", "local3", "local4") if l in spec.layers]:
            monitor.SummaryHist("weight", layer + "/weights:0", layer)
            monitor.SummaryHist("bias", layer + "/biases:0", layer)
            monitor.SummaryHist("activation", layer + "/" + layer + ":0", layer)
            monitor.SummaryNorm2("weight", layer + "/weights:0", layer)
        monitor.SummaryGradient("weight", "loss")
        monitor.SummaryGWRatio()
        monitor.SummaryScalar("train loss", "loss")
        monitor.SummaryScalar("train accuracy", "accuracy")
        monitor.SummaryScalar("test loss", "loss")
        monitor.SummaryScalar("test accuracy", "accuracy")
        hooks.append(trainer.LoggerHook(monitor, writer, test_interval))
        saver = checkpointing.CheckpointSaver(FLAGS.train_dir, spec, opt.kind, save_secs=hp.save_checkpoint_secs,
                                              num_shards=max(1, cluster.num_ps), max_to_keep=hp.max_to_keep)

    try:
        with trainer.MonitoredTrainingLoop(engine, source, hooks=hooks, is_chief=is_chief, saver=saver, task_id=role.task_id,
                                           feeder=feeder, verbose_steps=FLAGS.verbose_steps, group=group) as sess:
            while not sess.should_stop():
                sess.run_step()
    except BaseException:
        # leave the peers' mapped memory in an orderly way before the interpreter tears the CUDA context down
        if use_cuda:
            try:
                torch.cuda.synchronize()
            except Exception:
                pass
        if comm is not None:
            try:
                comm.close(collective=False)
            except Exception:
                pass
        raise
    m = engine.metrics()
    print("final: global_step=%d loss=%.4f accuracy=%.3f" % (m["global_step"], m["total_loss"], m["accuracy"]))

    if threaded:
        plane.barrier()
        comm.close()
    elif group is not None:
        import torch.distributed as dist

        cluster_spec.signal_worker_done(store)
        try:
            dist.barrier(group=group)
        except Exception:
            pass
        if comm is not None:
            comm.close()
        dist.destroy_process_group()
    if hasattr(source, "close"):
        source.close()
    return 0


def _strip(argv, names):
    """argv without the given flags ("--flag=v" and "--flag v" spellings)."""
    out, skip = [], False
    for a in argv:
        if skip:
            skip = False
        elif a.split("=", 1)[0] in names:
            skip = "=" not in a and a.split("=", 1)[0] not in ("--tower_procs",)
        else:
            out.append(a)
    return out


def main(argv=None):
    from distributed_tensorflow_ibm_mnist_b200.parallel import towers

    supervised = os.environ.get("DTM_SUPERVISED") == "1"
    k = FLAGS.num_gpus if (FLAGS.num_gpus and FLAGS.num_gpus > 1 and not FLAGS.worker_hosts) else 0
    use_threads = False
    if k and not FLAGS.tower_procs and FLAGS.engine != "torch":
        import torch

        use_threads = torch.cuda.is_available() and torch.cuda.device_count() >= k
    me = os.path.abspath(__file__)
    if not supervised:
        if k and not use_threads:       # one worker process per device
            return towers.launch_towers(k, _strip(sys.argv[1:], ("--num_gpus", "--max_restarts", "--tower_procs")), script=me,
                                        max_restarts=FLAGS.max_restarts)
        if FLAGS.max_restarts > 0 and not FLAGS.worker_hosts:      # one supervised process (which may itself drive k GPUs)
            return towers.launch_towers(1, _strip(sys.argv[1:], ("--max_restarts",)), script=me, max_restarts=FLAGS.max_restarts)
    if use_threads:
        hp = _configure_hyper_params()
        return towers.run_tower_threads(k, lambda role, plane: train(role, plane, hp))
    return train()


if __name__ == "__main__":
    flags.run(main)
