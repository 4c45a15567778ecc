"""The product path on a real GPU: the reference's entry points driving the sm_100a engine end to end.

* ``main.py --engine=native`` (local mode) -> TF-format checkpoint -> ``inference.py --engine=native`` (reference
  main.py:140-156 and inference.py:86-110), compared with an in-process restore of the same files;
* checkpoint round trip of the native engine: save -> restore into a fresh engine is bit exact for every saved variable
  (weights, optimizer slots, EMA shadows, step, beta powers, loss averages), and one more step from the restored state
  equals the uninterrupted run;
* two workers (``--worker_hosts``): fused peer-memory gradient exchange, sharded optimizer state gathered for the
  checkpoint (multi ``--ps_hosts`` shard layout), compared with a single-GPU run of the same global batch.  With two GPUs
  the ranks use one GPU each and NCCL as control plane; on a ONE-GPU box both ranks share the GPU (time-sliced contexts,
  CUDA-IPC peer pointers, gloo control plane) - slow, but it runs the same kernels and the same flag protocol;
* a vanished rank: the survivor's in-kernel watchdog must trap with a readable message within seconds, not hang;
* tower mode ``--num_gpus=2``; 300-step convergence of the bf16 engine against the fp32 PyTorch engine.
"""
import json
import os
import re
import socket
import subprocess
import sys
import time

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# a collective that cannot complete should fail these tests in seconds, not after the default two-minute leash
ENV = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2", DTM_COMM_WATCHDOG_MS=os.environ.get("DTM_COMM_WATCHDOG_MS", "20000"))


def _run(args, timeout=600, env=None):
    return subprocess.run([sys.executable, *args], cwd=ROOT, env=env or ENV, capture_output=True, text=True, timeout=timeout)


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _two_rank_env():
    """Two GPUs: one each, NCCL control plane.  One GPU: both ranks on it, gloo control plane (NCCL refuses duplicate GPUs)."""
    if torch.cuda.device_count() >= 2:
        return [dict(ENV, LOCAL_RANK=str(k)) for k in range(2)], "nccl"
    return [dict(ENV, LOCAL_RANK="0", DTM_NVLS="0") for _ in range(2)], "gloo"


def _write_pngs(d, n=7):
    from PIL import Image

    rng = np.random.default_rng(3)
    os.makedirs(d, exist_ok=True)
    imgs = rng.integers(0, 256, (n, 28, 28), dtype=np.uint8)
    for i in range(n):
        Image.fromarray(imgs[i], "L").save(os.path.join(d, f"img{i}.png"))
    return imgs


def test_native_main_checkpoint_resume_inference(tmp_path):
    train_dir = str(tmp_path / "train")
    common = ["main.py", f"--train_dir={train_dir}", "--model=baseline", "--batch_size=64", "--test_interval=5", "--optimizer=adam",
              "--base_lr=0.001", "--engine=native", "--save_checkpoint_secs=0", "--log_device_placement"]
    r = _run(common + ["--max_steps=12"])
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-2500:])
    assert "12 steps executed on worker 0." in r.stdout and "NativeEngine" in r.stdout
    assert os.path.exists(os.path.join(train_dir, "model.ckpt-12.index")) and os.path.exists(os.path.join(train_dir, "graph.pbtxt"))
    from distributed_tensorflow_ibm_mnist_b200.formats import events
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader
    import glob

    rd = BundleReader(os.path.join(train_dir, "model.ckpt-12"))
    assert int(rd.get("global_step")) == 12 and "local3/weights/Adam_1" in rd and "conv2/weights/ExponentialMovingAverage" in rd
    evs = events.read_events(glob.glob(os.path.join(train_dir, "log", "events.out.tfevents.*"))[0])
    tags = set().union(*[set(e["scalars"]) | set(e["histograms"]) for e in evs])
    assert {"train loss", "test loss", "test accuracy", "conv1/weight", "local3/gradient_weight_ratio"} <= tags
    # resume: the step budget is global
    r = _run(common + ["--max_steps=20"])
    assert r.returncode == 0, r.stderr[-2500:]
    assert "restored checkpoint at global step 12" in r.stdout and "8 steps executed on worker 0." in r.stdout

    # inference.py with the native engine on the same checkpoint directory
    img_dir, out_dir = str(tmp_path / "images"), str(tmp_path / "out")
    imgs = _write_pngs(img_dir)
    r = _run(["inference.py", f"--model={train_dir}", f"--input_dir={img_dir}", f"--output_dir={out_dir}", "--output_file=res.json",
              "--prob_thresh=0.0", "--engine=native"])
    assert r.returncode == 0, r.stderr[-2500:]
    doc = json.load(open(os.path.join(out_dir, "res.json")))
    assert doc["num_samples"] == 7
    # in-process restores of the same files: the native engine reproduces the script's probabilities exactly, the fp32
    # PyTorch engine within bf16 operand precision
    from distributed_tensorflow_ibm_mnist_b200.models import presets
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing as ck
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine
    from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

    spec = presets.make_spec("baseline")
    x = torch.from_numpy(imgs.reshape(7, 784))
    nat = NativeEngine(spec, 128, use_graph=False)
    assert ck.restore_checkpoint(train_dir, nat, spec, "sgd", model_only=True) == 20
    want_native = nat.predict(x.cuda()).cpu().numpy()
    ref = TorchEngine(spec, 128)
    ck.restore_checkpoint(train_dir, ref, spec, "sgd", model_only=True)
    want_fp32 = ref.predict(x).numpy()
    for k, res in enumerate(doc["results"]):
        got = {int(p["label"]): p["prob"] for p in res["predictions"]}
        row = np.array([got[c] for c in range(10)])
        assert np.allclose(row, want_native[k], atol=1e-6)
        assert np.abs(row - want_fp32[k]).max() < 2e-2


def _trained_engine(spec, B, steps, *, optimizer="adam", seed=5):
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.models import oracle
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    eng = NativeEngine(spec, B, optimizer=optimizer, base_lr=1e-3, seed=77, use_graph=True)
    eng.load_params(oracle.init_params(spec, seed=seed))
    src = SyntheticMnist(B, spec.image_pixels, pool_batches=4, seed=9, separable=True, pin=False)
    batches = [src.next() for _ in range(steps + 1)]
    for i in range(steps):
        eng.stage_batch(0, batches[i][0].cuda(), batches[i][1].cuda())
        eng.train_step(0)
    torch.cuda.synchronize()
    return eng, batches


@pytest.mark.parametrize("preset,optimizer", [("baseline", "adam"), ("reference", "momentum")])
def test_native_checkpoint_roundtrip_bit_exact_and_one_more_step(tmp_path, preset, optimizer):
    from distributed_tensorflow_ibm_mnist_b200.models import presets
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing as ck
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    spec = presets.make_spec(preset)
    B, K = 64, 6
    a, batches = _trained_engine(spec, B, K, optimizer=optimizer)
    ck.save_checkpoint(str(tmp_path), a, spec, optimizer, num_shards=2)
    saved = ck.collect_variables(a, spec, optimizer)
    b = NativeEngine(spec, B, optimizer=optimizer, base_lr=1e-3, seed=77, use_graph=True)
    assert ck.restore_checkpoint(str(tmp_path), b, spec, optimizer) == K
    restored = ck.collect_variables(b, spec, optimizer)
    assert set(saved) == set(restored)
    for k in saved:          # everything a TF Saver would hold, bit for bit
        assert np.array_equal(saved[k], restored[k]), k
    # one more step on the same batch from both: the uninterrupted engine and the restored one
    for eng in (a, b):
        eng.stage_batch(0, batches[K][0].cuda(), batches[K][1].cuda())
        eng.train_step(0)
    torch.cuda.synchronize()
    va, vb = ck.collect_variables(a, spec, optimizer), ck.collect_variables(b, spec, optimizer)
    assert int(va["global_step"]) == int(vb["global_step"]) == K + 1
    for k in va:
        if k.startswith("local3/weights") or k.startswith("local4/weights"):
            # gradients stored by a GEMM (no atomics): the restored run is bit identical to the uninterrupted one
            assert np.array_equal(va[k], vb[k]), k
        else:
            # conv / bias / head gradients are accumulated with floating-point atomics whose order differs from launch to
            # launch, so two runs of the SAME engine differ in the last bits as well
            assert np.allclose(va[k], vb[k], rtol=2e-4, atol=2e-6), (k, np.abs(va[k] - vb[k]).max())


def _spawn_two_workers(train_dir, extra, envs, backend, model="baseline"):
    ports = _free_ports(2)
    hosts = [f"--worker_hosts=127.0.0.1:{ports[0]},127.0.0.1:{ports[1]}"]     # no ps process: worker 0 hosts the rendezvous store
    args = ["main.py", f"--train_dir={train_dir}", f"--model={model}", "--engine=native", f"--backend={backend}", *hosts, *extra]
    return [subprocess.Popen([sys.executable, *args, "--job_name=worker", f"--task_id={k}"], cwd=ROOT, env=envs[k],
                             stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for k in range(2)]


def test_two_workers_fused_exchange_sharded_checkpoint(tmp_path):
    """The reference's own graph (``reference`` preset: no dropout, so the trajectory is deterministic up to summation order)
    trained by two workers must land where one GPU lands with the concatenated batch."""
    envs, backend = _two_rank_env()
    train_dir = str(tmp_path / "two")
    procs = _spawn_two_workers(train_dir, ["--batch_size=32", "--max_steps=6", "--test_interval=100", "--optimizer=momentum",
                                           "--base_lr=0.02", "--save_checkpoint_secs=0", "--seed=3"], envs, backend, model="reference")
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=600))
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    for p, (o, e) in zip(procs, outs):
        assert p.returncode == 0, (o[-1500:], e[-3000:])
    assert "6 steps executed on worker 0." in outs[0][0] and "6 steps executed on worker 1." in outs[1][0]
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader

    rd = BundleReader(os.path.join(train_dir, "model.ckpt-6"))
    assert int(rd.get("global_step")) == 6
    # the same job on one GPU with the global batch: both workers' batches concatenated (seed + rank), same initial weights
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    spec = presets.make_spec("reference")
    one = NativeEngine(spec, 64, optimizer="momentum", base_lr=0.02, decay_steps=presets.decay_steps(32), seed=3, use_graph=False)
    one.load_params(oracle.init_params(spec, seed=3))
    srcs = [SyntheticMnist(32, spec.image_pixels, pool_batches=16, seed=3 + r, separable=True, pin=False) for r in range(2)]
    for _ in range(6):
        parts = [s.next() for s in srcs]
        one.stage_batch(0, torch.cat([p[0] for p in parts]).cuda(), torch.cat([p[1] for p in parts]).cuda())
        one.train_step(0)
    torch.cuda.synchronize()
    want, want_ema, want_slot = one.params(), one.ema_params(), one.slots()[0]

    def rel(a, b):
        return np.abs(a - b).max() / (np.abs(b).max() + 1e-20)

    # weights are replicated; the momentum slots and EMA shadows of the fc matrices live on their owning rank only and reach
    # the file through the collective gather - a wrong gather leaves half of them at their initial values
    for name in ("conv1/weights", "conv2/weights", "local3/weights", "local4/weights", "softmax_linear/weights", "local3/biases"):
        assert rel(rd.get(name), want[name].numpy()) < 3e-3, (name, rel(rd.get(name), want[name].numpy()))
        # the slot is a sum of six gradients: rounding flips of bf16 activations (the two runs sum in different orders) show
        # up here first - judge it as a whole (relative L2) and bound the worst element loosely
        got_m, want_m = rd.get(name + "/Momentum"), want_slot[name].numpy()
        assert np.linalg.norm(got_m - want_m) / (np.linalg.norm(want_m) + 1e-20) < 4e-2, name
        assert rel(got_m, want_m) < 0.1, (name, rel(got_m, want_m))
        assert rel(rd.get(name + "/ExponentialMovingAverage"), want_ema[name].numpy()) < 3e-3, name
    v = rd.get("local3/weights/Momentum").reshape(-1)
    half = v.size // 2
    assert np.count_nonzero(v[:half]) > 0.1 * half and np.count_nonzero(v[half:]) > 0.1 * half


def test_missing_rank_trips_the_kernel_watchdog(tmp_path):
    """Worker 1 dies without a goodbye after 3 steps; worker 0's next exchange kernel waits for its flag, the watchdog (set to
    3 s here) prints what it waited for and traps, and the process ends with an error instead of hanging the GPU."""
    envs, backend = _two_rank_env()
    envs = [dict(e, DTM_COMM_WATCHDOG_MS="3000", DTM_FAULT_INJECT="1:3") for e in envs]
    t0 = time.time()
    procs = _spawn_two_workers(str(tmp_path / "gone"), ["--batch_size=32", "--max_steps=400", "--save_checkpoint_secs=-1"], envs, backend)
    try:
        o1, e1 = procs[1].communicate(timeout=300)
        o0, e0 = procs[0].communicate(timeout=120)
    except subprocess.TimeoutExpired:
        for q in procs:
            q.kill()
        raise
    assert procs[1].returncode == 43 and "[fault-injection] worker 1 dies" in o1
    assert procs[0].returncode != 0
    assert "[dtm watchdog] peer barrier (a rank did not arrive)" in (o0 + e0), (o0[-1500:], e0[-1500:])
    assert time.time() - t0 < 280


def test_tower_mode_on_gpus(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("tower mode places one worker per GPU: needs 2 GPUs")
    train_dir = str(tmp_path / "towers")
    r = _run(["main.py", "--num_gpus=2", f"--train_dir={train_dir}", "--model=baseline", "--batch_size=64", "--max_steps=10",
              "--test_interval=4", "--engine=native", "--save_checkpoint_secs=-1"])
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-2500:])
    assert "10 steps executed on worker 0." in r.stdout and "10 steps executed on worker 1." in r.stdout
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader

    assert int(BundleReader(os.path.join(train_dir, "model.ckpt-10")).get("global_step")) == 10


def test_single_process_towers_match_process_towers(tmp_path):
    """``--num_gpus=2``: one process, one thread + engine per GPU, symmetric memory from our VMM allocator (no IPC, no ports, no
    torch.distributed) - must land where the two-process form (``--tower_procs``) lands."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader

    common = ["main.py", "--num_gpus=2", "--model=reference", "--batch_size=32", "--max_steps=8", "--test_interval=100",
              "--optimizer=momentum", "--base_lr=0.02", "--engine=native", "--save_checkpoint_secs=-1", "--seed=5"]
    r1 = _run(common + [f"--train_dir={tmp_path / 'threads'}"])
    assert r1.returncode == 0, (r1.stdout[-1500:], r1.stderr[-3000:])
    assert "single-process tower mode: 2 GPUs driven by pid" in r1.stdout and "[supervisor]" not in r1.stdout
    assert "8 steps executed on worker 0." in r1.stdout and "8 steps executed on worker 1." in r1.stdout
    r2 = _run(common + [f"--train_dir={tmp_path / 'procs'}", "--tower_procs"])
    assert r2.returncode == 0, (r2.stdout[-1500:], r2.stderr[-3000:])
    a, b = BundleReader(str(tmp_path / "threads" / "model.ckpt-8")), BundleReader(str(tmp_path / "procs" / "model.ckpt-8"))
    assert int(a.get("global_step")) == int(b.get("global_step")) == 8
    exact = 0
    for name in ("conv1/weights", "conv2/weights", "local3/weights", "local4/weights", "softmax_linear/weights",
                 "local3/weights/Momentum", "local3/weights/ExponentialMovingAverage"):
        x, y = a.get(name), b.get(name)
        exact += int(np.array_equal(x, y))
        # floating-point atomics reorder sums, and a flipped bf16 rounding downstream of them shows up first in the slot (a sum
        # of gradients): weights and shadows element-wise, the slot as a whole
        if name.endswith("/Momentum"):
            assert np.linalg.norm(x - y) / (np.linalg.norm(y) + 1e-20) < 4e-2 and np.abs(x - y).max() <= 0.1 * np.abs(y).max(), name
        else:
            assert np.abs(x - y).max() <= 2e-3 * (np.abs(y).max() + 1e-20), name
    print("bit-identical variables:", exact, "of 7")


def test_convergence_bf16_engine_tracks_fp32_engine():
    """300 steps on separable synthetic data: the bf16 tensor-core engine and the fp32 PyTorch engine (the reference's
    precision, mnist_input.py:39/:75-88) start from the same weights, see the same batches and the same dropout masks.  The
    smoothed loss curves must stay within a band and both must learn the task."""
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine
    from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

    spec = presets.make_spec("baseline")
    B, steps = 128, 300
    kw = dict(optimizer="adam", base_lr=1e-3, decay_steps=presets.decay_steps(B), seed=21)
    nat = NativeEngine(spec, B, use_graph=True, **kw)
    ref = TorchEngine(spec, B, device="cuda", **kw)
    init = oracle.init_params(spec, seed=4)
    nat.load_params(init)
    ref.load_params(init)
    src = SyntheticMnist(B, spec.image_pixels, pool_batches=32, seed=11, separable=True, pin=False)
    ln, lr_, an, ar = [], [], [], []
    for _ in range(steps):
        x, y = src.next()
        nat.stage_batch(0, x.cuda(), y.cuda())
        nat.train_step(0)
        ref.stage_batch(0, x, y)
        ref.train_step(0)
        m, r = nat.metrics(), ref.metrics()
        ln.append(m["cross_entropy"]); lr_.append(r["cross_entropy"]); an.append(m["accuracy"]); ar.append(r["accuracy"])
    ln, lr_ = np.array(ln), np.array(lr_)
    assert not nat.metrics()["nan"]
    smooth = lambda v: np.convolve(v, np.ones(20) / 20, mode="valid")
    sn, sr = smooth(ln), smooth(lr_)
    assert np.abs(sn - sr).max() < 0.08 * max(1.0, sr.max()), (np.abs(sn - sr).max(), sn[-1], sr[-1])
    assert sn[-1] < 0.25 * sn[0] and sr[-1] < 0.25 * sr[0]            # both learned
    assert np.mean(an[-20:]) > 0.9 and abs(np.mean(an[-20:]) - np.mean(ar[-20:])) < 0.03


def test_tfrecord_training_through_feeder_and_batch_ring(tmp_path):
    """Real-data input on the GPU: TFRecords -> C++ loader (batch ring of pinned buffers) -> DeviceFeeder's copy stream ->
    device slots.  (a) With a slow consumer and the smallest ring every lent buffer is recycled while copies are still
    queued: each step must still see exactly its batch.  (b) The native engine trains from the same files."""
    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import DeviceFeeder, SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord
    from distributed_tensorflow_ibm_mnist_b200.models import presets
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    B, n = 64, 64 * 24
    src = SyntheticMnist(B, 784, pool_batches=24, seed=2, separable=True, pin=False)
    imgs, labels = [], []
    for _ in range(24):
        x, y = src.next()
        imgs += [bytes(r.numpy().tobytes()) for r in x.view(B, 784)]
        labels += [int(v) for v in y]
    path = str(tmp_path / "train.tfrecord")
    assert tfrecord.write_mnist_tfrecord(path, imgs, labels) == n

    class Probe:                                  # an engine that only records what each step was given
        def __init__(self):
            self.B, self.device = B, torch.device("cuda", 0)
            self.spec = presets.make_spec("baseline")
            self.packed_label_offset = B * 784
            self.xy = [torch.zeros(B * 784 + 4 * B, dtype=torch.uint8, device="cuda") for _ in range(2)]
            self.seen = []

        def stage_packed(self, slot, packed, stream=None):
            with torch.cuda.stream(stream):
                self.xy[slot].copy_(packed, non_blocking=True)

        def train_step(self, slot):
            torch.cuda._sleep(400_000)           # ~0.2 ms: the host runs far ahead of the GPU
            self.seen.append(self.xy[slot].clone())

    probe = Probe()
    ds = readers.TFRecordBatches([path], B, 784, shuffle=False, native=True, pin=True, prefetch_batches=2)
    assert isinstance(ds, readers.NativeRecordBatches) and ds.label_offset == probe.packed_label_offset
    feeder = DeviceFeeder(probe, ds)
    assert feeder._packed
    for _ in range(60):
        feeder.step()
    torch.cuda.synchronize()
    for k, got in enumerate(probe.seen):
        lo = (k * B) % n
        want = b"".join(imgs[lo:lo + B])
        assert bytes(got[:B * 784].cpu().numpy().tobytes()) == want, k
        assert [int(v) for v in got[B * 784:].view(torch.int32).cpu()] == labels[lo:lo + B], k
    ds.close()

    from distributed_tensorflow_ibm_mnist_b200.models import oracle

    eng = NativeEngine(presets.make_spec("baseline"), B, optimizer="adam", base_lr=1e-3, seed=3, use_graph=True)
    eng.load_params(oracle.init_params(eng.spec, seed=4))
    ds = readers.TFRecordBatches([path], B, 784, seed=4, min_after_dequeue=256, native=True, pin=True)
    feeder = DeviceFeeder(eng, ds)
    first = None
    for i in range(150):
        feeder.step()
        if i == 4:
            first = eng.metrics()["cross_entropy"]
    last = eng.metrics()["cross_entropy"]
    ds.close()
    assert np.isfinite(first) and last < 0.7 * first, (first, last)
