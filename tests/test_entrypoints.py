"""End-to-end plumbing on CPU (BASELINE.json configs #1 and #2): the three entry points as real processes.

* local mode: ``main.py`` trains, writes TF-format checkpoints + event files, resumes; ``inference.py`` restores them.
* between-graph replication: 1 ps + 2 workers on one host over gloo; the ps process returns once the workers finish;
  ``inference.py`` on the resulting checkpoint reproduces the logits of an in-process restore bit for bit.
"""
import glob
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=ROOT, OMP_NUM_THREADS="2")


def _run(args, timeout=240, **kw):
    return subprocess.run([sys.executable, *args], cwd=ROOT, env=ENV, capture_output=True, text=True, timeout=timeout, **kw)


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _write_pngs(d, n=5):
    from PIL import Image

    rng = np.random.default_rng(0)
    os.makedirs(d, exist_ok=True)
    imgs = rng.integers(0, 256, (n, 28, 28), dtype=np.uint8)
    for i in range(n):
        Image.fromarray(imgs[i], "L").save(os.path.join(d, f"img{i}.png"))
    open(os.path.join(d, "notes.txt"), "w").write("ignored")
    return imgs


def test_local_training_checkpoint_resume_and_inference(tmp_path):
    train_dir = str(tmp_path / "train")
    common = ["main.py", f"--train_dir={train_dir}", "--model=reference", "--batch_size=16", "--test_interval=2",
              "--optimizer=momentum", "--base_lr=0.01", "--engine=torch", "--save_checkpoint_secs=0", "--log_device_placement"]
    r = _run(common + ["--max_steps=5"])
    assert r.returncode == 0, r.stderr[-2000:]
    assert "5 steps executed on worker 0." in r.stdout and "[placement]" in r.stdout
    assert os.path.exists(os.path.join(train_dir, "model.ckpt-5.index"))
    assert os.path.exists(os.path.join(train_dir, "model.ckpt-5.data-00000-of-00001"))
    # the chief's MonitoredTrainingSession artefacts: graph.pbtxt, a MetaGraphDef per checkpoint, an event file in train_dir
    assert os.path.exists(os.path.join(train_dir, "graph.pbtxt")) and os.path.exists(os.path.join(train_dir, "model.ckpt-5.meta"))
    assert 'name: "local4/weights/Momentum"' in open(os.path.join(train_dir, "graph.pbtxt")).read()
    assert len(glob.glob(os.path.join(train_dir, "events.out.tfevents.*"))) == 1
    from distributed_tensorflow_ibm_mnist_b200.formats import checkpoint_state as cs, events
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader

    assert cs.read_state(train_dir).model_checkpoint_path == "model.ckpt-5"
    rd = BundleReader(os.path.join(train_dir, "model.ckpt-5"))
    assert int(rd.get("global_step")) == 5 and rd.get("conv1/weights").shape == (5, 5, 3, 32)
    assert "local3/weights/Momentum" in rd and "local4/weights/ExponentialMovingAverage" in rd
    # chief summaries: steps at multiples of test_interval fetch train + test collections (main.py:125-135)
    ev_files = glob.glob(os.path.join(train_dir, "log", "events.out.tfevents.*"))
    assert len(ev_files) == 1
    evs = events.read_events(ev_files[0])
    tags = set().union(*[set(e["scalars"]) | set(e["histograms"]) for e in evs])
    assert {"train loss", "train accuracy", "test loss", "test accuracy", "conv1/weight", "local4/activation",
            "local3/weight_norm2", "conv2/weight_gradient", "local3/gradient_weight_ratio"} <= tags
    assert sorted({e["step"] for e in evs if e["scalars"]}) == [3, 5]
    # resume: global step budget is global (StopAtStepHook(last_step)), so only 3 more steps run
    r = _run(common + ["--max_steps=8"])
    assert r.returncode == 0, r.stderr[-2000:]
    assert "restored checkpoint at global step 5" in r.stdout and "3 steps executed on worker 0." in r.stdout
    assert os.path.exists(os.path.join(train_dir, "model.ckpt-8.index"))

    # inference.py, image-directory mode (1-channel PNGs into the 3-channel checkpoint: quirk B1 handled)
    img_dir = str(tmp_path / "images")
    imgs = _write_pngs(img_dir)
    out_dir = str(tmp_path / "out")
    r = _run(["inference.py", f"--model={train_dir}", f"--input_dir={img_dir}", f"--output_dir={out_dir}", "--output_file=res.json",
              "--prob_thresh=0.0", "--engine=torch"])
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Predictions are Finished in" in r.stdout
    doc = json.load(open(os.path.join(out_dir, "res.json")))
    assert doc["num_samples"] == 5 and [os.path.basename(x["sampleId"]) for x in doc["results"]] == [f"img{i}.png" for i in range(5)]
    # identical probabilities from an in-process restore of the same checkpoint
    from distributed_tensorflow_ibm_mnist_b200.models import presets
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing as ck
    from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

    spec = presets.make_spec("reference")
    eng = TorchEngine(spec, 128)
    ck.restore_checkpoint(train_dir, eng, spec, "sgd", model_only=True)
    x = torch.from_numpy(np.repeat(imgs.reshape(5, 784, 1), 3, axis=2).reshape(5, 2352))
    want = eng.predict(x).numpy()
    for row, res in zip(want, doc["results"]):
        got = {int(p["label"]): p["prob"] for p in res["predictions"]}
        assert np.allclose([got[c] for c in range(10)], row, atol=1e-6)
    # no checkpoint: say so, predict nothing, still write the (empty) result
    r = _run(["inference.py", f"--model={tmp_path / 'void'}", f"--input_dir={img_dir}", f"--output_dir={out_dir}", "--output_file=none.json"])
    assert r.returncode == 0 and "no checkpoint found" in r.stderr
    assert json.load(open(os.path.join(out_dir, "none.json")))["num_samples"] == 0


def test_ps_plus_two_workers_gloo_and_validation_inference(tmp_path):
    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord

    train_dir = str(tmp_path / "train")
    p_ps, p_w0, p_w1 = _free_ports(3)
    hosts = [f"--ps_hosts=127.0.0.1:{p_ps}", f"--worker_hosts=127.0.0.1:{p_w0},127.0.0.1:{p_w1}"]
    common = ["main.py", f"--train_dir={train_dir}", "--model=baseline", "--batch_size=8", "--max_steps=4", "--test_interval=2",
              "--optimizer=adam", "--base_lr=0.001", "--engine=torch", "--save_checkpoint_secs=0", *hosts]
    procs = [subprocess.Popen([sys.executable, *common, "--job_name=ps", "--task_index=0"], cwd=ROOT, env=ENV,
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)]
    for k in range(2):
        procs.append(subprocess.Popen([sys.executable, *common, "--job_name=worker", f"--task_id={k}"], cwd=ROOT, env=ENV,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=300)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append((p.returncode, o, e))
    for rc, o, e in outs:
        assert rc == 0, e[-2000:]
    assert "Cluster spec: " in outs[1][1]
    assert "4 steps executed on worker 0." in outs[1][1] and "4 steps executed on worker 1." in outs[2][1]
    # synchronous data parallelism: both workers end at the same global step with the same loss trajectory end point
    assert os.path.exists(os.path.join(train_dir, "model.ckpt-4.index"))
    assert glob.glob(os.path.join(train_dir, "log", "events.out.tfevents.*"))          # chief only

    # inference.py --validate on TFRecords (784-byte records, as inference.py:59 assumes)
    rng = np.random.default_rng(1)
    imgs = rng.integers(0, 256, (6, 784), dtype=np.uint8)
    labels = [int(v) for v in rng.integers(0, 10, 6)]
    rec = str(tmp_path / "val.tfrecord")
    tfrecord.write_mnist_tfrecord(rec, [i.tobytes() for i in imgs], labels)
    out_dir = str(tmp_path / "out")
    r = _run(["inference.py", f"--model={train_dir}", "--validate", f"--val_data={rec}", f"--output_dir={out_dir}",
              "--output_file=val.json", "--engine=torch"])
    assert r.returncode == 0, r.stderr[-2000:]
    doc = json.load(open(os.path.join(out_dir, "val.json")))
    assert doc["num_samples"] == 6 and [int(x["groundTruth"]) for x in doc["results"]] == labels
    from distributed_tensorflow_ibm_mnist_b200.models import presets
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing as ck
    from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

    spec = presets.make_spec("baseline")
    eng = TorchEngine(spec, 128)
    ck.restore_checkpoint(train_dir, eng, spec, "sgd", model_only=True)
    want = eng.predict(torch.from_numpy(imgs)).numpy()
    for row, res in zip(want, doc["results"]):
        assert int(res["label"]) == int(np.argmax(row)) and abs(res["prob"] - row.max()) < 1e-6
    assert doc["accuracy"] == float(np.mean([int(np.argmax(r)) == l for r, l in zip(want, labels)]))


def test_tower_mode_spawns_one_worker_per_device(tmp_path):
    """``--num_gpus=k`` (SURVEY D5 / P-i): one worker process per device on this node, synchronous steps, the chief writes
    the checkpoints.  On this CPU box the towers fall back to gloo + the portable engine."""
    train_dir = str(tmp_path / "towers")
    r = _run(["main.py", "--num_gpus=2", f"--train_dir={train_dir}", "--model=baseline", "--batch_size=8", "--max_steps=6",
              "--test_interval=3", "--optimizer=momentum", "--base_lr=0.02", "--engine=torch", "--save_checkpoint_secs=-1"])
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert "6 steps executed on worker 0." in r.stdout and "6 steps executed on worker 1." in r.stdout
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader

    rd = BundleReader(os.path.join(train_dir, "model.ckpt-6"))
    assert int(rd.get("global_step")) == 6 and rd.get("conv1/weights").shape == (5, 5, 1, 32)
    assert os.path.exists(os.path.join(train_dir, "graph.pbtxt"))


def test_supervisor_restarts_failed_worker_set_from_checkpoint(tmp_path):
    """Recoverable session (reference main.py:140-146 -> _RecoverableSession): worker 1 of a two-tower job
    dies without clean-up after 4 steps; the supervisor stops the survivor, restarts both, they restore the chief's newest
    checkpoint and finish the GLOBAL step budget."""
    train_dir = str(tmp_path / "recover")
    env = dict(ENV, DTM_FAULT_INJECT="1:4")
    r = subprocess.run([sys.executable, "main.py", "--num_gpus", "2", "--max_restarts=1", f"--train_dir={train_dir}", "--model=baseline",
                        "--batch_size=8", "--max_steps=9", "--test_interval=100", "--optimizer=momentum", "--base_lr=0.02",
                        "--engine=torch", "--save_checkpoint_secs=0"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=400)
    assert r.returncode == 0, (r.stdout[-2500:], r.stderr[-2500:])
    assert "[fault-injection] worker 1 dies after local step 4" in r.stdout
    assert "[supervisor] worker 1 exited with status 43; restarting all 2 worker(s) from the newest checkpoint (restart 1 of 1)" in r.stdout
    import re

    m = re.search(r"restored checkpoint at global step (\d+)", r.stdout)
    assert m and 1 <= int(m.group(1)) <= 5, r.stdout[-2000:]
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader

    assert int(BundleReader(os.path.join(train_dir, "model.ckpt-9")).get("global_step")) == 9
    # without restarts left the failure is the job's result
    r = subprocess.run([sys.executable, "main.py", "--num_gpus=2", f"--train_dir={tmp_path / 'norestart'}", "--model=baseline",
                        "--batch_size=8", "--max_steps=9", "--engine=torch", "--save_checkpoint_secs=-1"], cwd=ROOT,
                       env=dict(ENV, DTM_FAULT_INJECT="0:2"), capture_output=True, text=True, timeout=400)
    assert r.returncode == 43 and "no restarts left" in r.stdout


def test_training_from_tfrecord_files_and_loader_bench(tmp_path):
    """Real-data route of main.py (mnist_input.inputs): --train_data TFRecords -> C++ loader (reader threads, shuffle queue,
    batch ring) -> engine; and the loader benchmark tool runs on a GPU-less host."""
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord

    src = SyntheticMnist(32, 784, pool_batches=20, seed=1, separable=True, pin=False)
    imgs, labels = [], []
    for _ in range(20):
        x, y = src.next()
        imgs += [bytes(r.numpy().tobytes()) for r in x.view(32, 784)]
        labels += [int(v) for v in y]
    files = [str(tmp_path / "a.tfrecords"), str(tmp_path / "b.tfrecords")]
    tfrecord.write_mnist_tfrecord(files[0], imgs[:320], labels[:320])
    tfrecord.write_mnist_tfrecord(files[1], imgs[320:], labels[320:])
    train_dir = str(tmp_path / "train")
    r = _run(["main.py", f"--train_dir={train_dir}", "--model=baseline", "--batch_size=32", "--test_interval=10", "--optimizer=adam",
              "--base_lr=0.001", "--engine=torch", "--save_checkpoint_secs=0", "--max_steps=45", "--input_threads=3",
              "--train_data=" + ",".join(files)])
    assert r.returncode == 0, r.stderr[-2500:]
    assert "45 steps executed on worker 0." in r.stdout and "image Shape" in r.stdout
    from distributed_tensorflow_ibm_mnist_b200.formats import events

    evs = events.read_events(glob.glob(os.path.join(train_dir, "log", "events.out.tfevents.*"))[0])
    losses = [e["scalars"]["train loss"] for e in evs if "train loss" in e["scalars"]]
    assert len(losses) == 4 and losses[-1] < losses[0]                 # separable classes: the loss falls within 40 steps
    r = _run(["main.py", f"--train_dir={train_dir}2", "--engine=torch", "--max_steps=2", "--train_data=" + str(tmp_path / "nope.tfrecords")])
    assert r.returncode != 0 and "Failed to find file" in (r.stderr + r.stdout)     # mnist_input.py:55
    r = _run(["tools/loader_bench.py", "--records", "3000", "--seconds", "0.5", "--threads", "2", "--prefetch", "3", "0"])
    assert r.returncode == 0, r.stderr[-1500:]
    rows = [json.loads(line) for line in r.stdout.strip().splitlines()]
    assert [row["prefetch_batches"] for row in rows] == [3, 0] and all(row["native"] and row["images_per_s"] > 0 for row in rows)
