"""CPU tests of the framework layer: flags, parameter manager, cluster roles, model semantics (closed-form checks from
SURVEY §2f), optimizer / EMA formulas, checkpoint save-resume, data readers, dataset conversion, monitor."""
import json
import math
import os
import struct
import sys

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from distributed_tensorflow_ibm_mnist_b200.cluster import spec as cluster_spec
from distributed_tensorflow_ibm_mnist_b200.config import flags as flags_mod
from distributed_tensorflow_ibm_mnist_b200.config import parameter_mgr as pm
from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- flags ------------------------------------------------------------------------------------------------
def _fresh_flags():
    fv = flags_mod._FlagValues()
    fv._define("job_name", "", "", str)
    fv._define("task_id", 0, "", int)
    fv._define("lr", 0.5, "", float)
    fv._define("log_device_placement", False, "", flags_mod._to_bool)
    fv.alias("task_index", "task_id")
    return fv


def test_flags_tf_style_syntax():
    fv = _fresh_flags()
    rest = fv.parse(["--job_name=worker", "--task_index", "3", "-lr=0.25", "--log_device_placement", "positional"])
    assert (fv.job_name, fv.task_id, fv.task_index, fv.lr, fv.log_device_placement) == ("worker", 3, 3, 0.25, True)
    assert rest == ["positional"]
    fv.parse(["--nolog_device_placement"])
    assert fv.log_device_placement is False
    fv.parse(["--log_device_placement=true", "--task_id=7"])
    assert fv.log_device_placement is True and fv.task_id == 7
    with pytest.raises(flags_mod.FlagError):
        fv.parse(["--unknown=1"])
    with pytest.raises(flags_mod.FlagError):
        fv.parse(["--task_id=abc"])
    assert fv.parse(["--unknown=1"], known_only=True) == ["--unknown=1"]
    fv.task_id = 0      # the reference assigns FLAGS.task_id = 0 in local mode (main.py:66)
    assert fv.task_id == 0


def test_entry_points_print_their_flags_on_help():
    """tf.app.run's --help: the program's docstring and flags, exit status 0; an unknown flag is still an error (status 2)."""
    import subprocess

    for script, flag in (("main.py", "--input_threads"), ("inference.py", "--prob_thresh")):
        for h in ("--help", "-h"):
            r = subprocess.run([sys.executable, script, h], cwd=ROOT, capture_output=True, text=True, timeout=120)
            assert r.returncode == 0 and "flags:" in r.stdout and flag in r.stdout and "flag error" not in r.stderr, (script, h)
    r = subprocess.run([sys.executable, "main.py", "--no_such_flag=1"], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 2 and "unknown flag --no_such_flag" in r.stderr


# ---- parameter manager -------------------------------------------------------------------------------------
def test_parameter_manager_getters_and_file(tmp_path):
    pm.reset()
    assert pm.getTrainBatchSize() == 128 and pm.getMaxSteps() == 1000
    f = tmp_path / "hp.json"
    f.write_text(json.dumps({"max_steps": 42, "test_interval": 7, "train_batch_size": 64, "base_learning_rate": 0.05,
                             "learning_rate_decay": 0.5, "optimizer": "momentum", "train_data": "a.tfr,b.tfr"}))
    pm.load_file(str(f))
    assert (pm.getMaxSteps(), pm.getTestInterval(), pm.getTrainBatchSize()) == (42, 7, 64)
    assert (pm.getBaseLearningRate(), pm.getLearningRateDecay()) == (0.05, 0.5)
    assert pm.getTrainData() == ["a.tfr", "b.tfr"] and pm.getTestData() == [] and pm.getValData() == []
    opt = pm.getOptimizer(0.05)
    assert opt.kind == "momentum" and opt.slot_names() == ["Momentum"]
    with pytest.raises(KeyError):
        pm.configure(not_a_param=1)
    pm.configure(optimizer="rmsprop")
    with pytest.raises(ValueError):
        pm.getOptimizer(0.1)
    pm.reset()


# ---- cluster ---------------------------------------------------------------------------------------------
def test_cluster_roles_follow_reference_rules():
    c = cluster_spec.ClusterSpec.from_flags("h:1", "h:2,h:3")
    assert c.as_dict() == {"worker": ["h:2", "h:3"], "ps": ["h:1"]} and not c.is_local
    assert c.rendezvous() == ("h", 1)
    r = cluster_spec.resolve_role(c, "worker", 1)
    assert (r.rank, r.world_size, r.is_chief) == (1, 2, False)
    assert cluster_spec.resolve_role(c, "worker", 0).is_chief
    assert cluster_spec.resolve_role(c, "ps", 0).rank == -1
    with pytest.raises(ValueError):
        cluster_spec.resolve_role(c, "worker", 2)
    with pytest.raises(ValueError):
        cluster_spec.resolve_role(c, "", 0)
    # ps hosts alone do not make a cluster: local mode, task 0, chief (main.py:56, :63-66)
    loc = cluster_spec.ClusterSpec.from_flags("h:1", "")
    assert loc.is_local
    r = cluster_spec.resolve_role(loc, "worker", 5)
    assert (r.task_id, r.is_chief, r.world_size) == (0, True, 1)
    assert cluster_spec.ClusterSpec.from_flags("", "h:9").rendezvous() == ("h", 9)


# ---- model semantics -----------------------------------------------------------------------------------------
def test_parameter_counts_and_schedule_constants():
    assert presets.make_spec("reference").num_params == 3_464_714
    assert presets.make_spec("reference", channels=1).num_params == 3_463_114
    assert presets.make_spec("baseline").num_params == 3_274_634
    assert presets.make_spec("wide").num_params == 25_832_330
    assert presets.decay_steps(128) == 136_718
    names = [v.name for v in presets.make_spec("reference").variables]
    assert names == ["conv1/weights", "conv1/biases", "conv2/weights", "conv2/biases", "local3/weights", "local3/biases",
                     "local4/weights", "local4/biases", "softmax_linear/weights", "softmax_linear/biases"]
    assert presets.make_spec("reference").var("conv1/weights").shape == (5, 5, 3, 32)


def test_initialisers_match_reference():
    spec = presets.make_spec("reference")
    p = oracle.init_params(spec, seed=1)
    assert torch.all(p["conv1/biases"] == 0) and torch.all(p["conv2/biases"] == 0.1) and torch.all(p["local3/biases"] == 0.1)
    w = p["local3/weights"]
    assert w.abs().max() <= 2 * 0.04 + 1e-7 and abs(w.std().item() - 0.04 * 0.8796) < 2e-3   # truncated at 2 sigma
    assert p["softmax_linear/weights"].abs().max() <= 2 / 192 + 1e-9


def test_tf_lrn_semantics():
    x = torch.randn(2, 12, 3, 3)
    got = oracle.tf_lrn(x)
    want = torch.empty_like(x)
    for c in range(12):
        lo, hi = max(0, c - 4), min(12, c + 5)
        s = (x[:, lo:hi] ** 2).sum(1)
        want[:, c] = x[:, c] / (1.0 + (0.001 / 9.0) * s) ** 0.75
    assert torch.allclose(got, want, atol=1e-6)


def test_lrn_kernel_arithmetic_in_float32():
    """The LRN epilogues of the kernels (csrc/common/ptx.cuh: window9_sums, lrn_pow_m075) restated in numpy float32: channel
    windows as differences of running prefix sums, s^-0.75 as r*sqrt(r) with r = s^-0.5.  Against float64 TF semantics the
    forward value must hold to 1e-5 for ordinary activations and stay bounded by alpha * ulp(prefix) when one channel dwarfs
    the others (the prefix then swallows the small squares - the error enters s = 1 + alpha*sum scaled by alpha = 1.1e-4)."""
    alpha, beta = np.float32(0.001 / 9.0), 0.75

    def lrn_kernel_way(a):                       # a: [pixels, C] float32
        n, c = a.shape
        ext = np.zeros((n, c + 8), np.float32)
        ext[:, 4:4 + c] = a * a
        p = np.zeros((n, c + 9), np.float32)
        for k in range(c + 8):                   # sequential float32 additions, like the unrolled loop
            p[:, k + 1] = p[:, k] + ext[:, k]
        win = p[:, 9:9 + c] - p[:, 0:c]
        s = np.float32(1.0) + alpha * win
        r = (np.float32(1.0) / np.sqrt(s)).astype(np.float32)
        return (a * (r * np.sqrt(r))).astype(np.float32)

    def lrn_reference(a):
        a64 = a.astype(np.float64)
        out = np.empty_like(a64)
        for ch in range(a.shape[1]):
            lo, hi = max(0, ch - 4), min(a.shape[1], ch + 5)
            out[:, ch] = a64[:, ch] / (1.0 + (0.001 / 9.0) * (a64[:, lo:hi] ** 2).sum(1)) ** beta
        return out

    rng = np.random.default_rng(0)
    for scale in (0.1, 1.0, 30.0):
        a = np.maximum(rng.normal(0, scale, (500, 64)), 0).astype(np.float32)
        got, want = lrn_kernel_way(a), lrn_reference(a)
        assert np.abs(got - want).max() <= 1e-5 * max(1.0, np.abs(want).max()), scale
    a = np.full((4, 32), 0.01, np.float32)
    a[:, 3] = 3000.0                             # prefix ~9e6: its ulp (1.0) swallows the 1e-4 squares that follow
    got, want = lrn_kernel_way(a), lrn_reference(a)
    far = slice(12, 32)                          # channels whose window does not contain the large one
    assert np.abs(got[:, far] / want[:, far] - 1).max() < 2e-4
    assert np.abs(got[:, :12] / want[:, :12] - 1).max() < 1e-5


def test_loss_is_xent_plus_weight_decay_and_accuracy_rule():
    spec = presets.make_spec("reference")
    p = oracle.init_params(spec, seed=0)
    x = torch.rand(8, spec.image_pixels) - 0.5
    y = torch.randint(0, 10, (8,))
    logits = oracle.forward(p, x, spec)
    assert logits.shape == (8, 10)
    total, xent = oracle.loss_fn(logits, y, p, spec)
    wd = 0.004 * 0.5 * (p["local3/weights"] ** 2).sum() + 0.004 * 0.5 * (p["local4/weights"] ** 2).sum()
    assert abs(float(total) - float(xent + wd)) < 1e-5
    assert abs(float(xent) - float(F.cross_entropy(logits, y))) < 1e-6
    lg = torch.tensor([[1.0, 1.0, 0.0], [0.0, 2.0, 1.0]])
    assert float(oracle.accuracy_fn(lg, torch.tensor([0, 2]))) == 0.5   # tie counts as correct, lower score does not


def test_forward_shapes_all_presets():
    for name in ("reference", "baseline", "wide"):
        spec = presets.make_spec(name)
        p = oracle.init_params(spec, seed=0)
        logits, acts = oracle.forward(p, torch.zeros(2, spec.image_pixels), spec, return_activations=True)
        assert logits.shape == (2, 10) and acts["conv1"].shape == (2, 32, 28, 28) and acts["conv2"].shape == (2, 64, 14, 14)


def test_learning_rate_and_ema_formulas():
    assert oracle.learning_rate(0.1, 0.5, 10, 9) == 0.1
    assert oracle.learning_rate(0.1, 0.5, 10, 10) == 0.05
    assert oracle.learning_rate(0.1, 0.5, 10, 25) == 0.025
    assert oracle.ema_keep(0) == 0.1 and abs(oracle.ema_keep(90) - 0.91) < 1e-12 and oracle.ema_keep(10 ** 7) == 0.9999


@pytest.mark.parametrize("kind", ["sgd", "momentum", "adam"])
def test_reference_optimizer_matches_hand_computation(kind):
    spec = presets.make_spec("baseline")
    w0 = torch.tensor([1.0, -2.0])
    g = torch.tensor([0.5, 0.25])
    params = {"w": w0.clone()}
    opt = oracle.ReferenceOptimizer(params, spec, optimizer=kind, base_lr=0.1, decay_rate=0.5, decay_steps=1, momentum=0.9)
    opt.step({"w": g})
    opt.step({"w": g})
    if kind == "sgd":
        want = w0 - 0.1 * g - 0.05 * g
    elif kind == "momentum":
        a1 = g
        a2 = 0.9 * a1 + g
        want = w0 - 0.1 * a1 - 0.05 * a2
    else:
        m1, v1 = 0.1 * g, 0.001 * g * g
        s1 = w0 - 0.1 * math.sqrt(1 - 0.999) / (1 - 0.9) * m1 / (v1.sqrt() + 1e-8)
        m2, v2 = 0.9 * m1 + 0.1 * g, 0.999 * v1 + 0.001 * g * g
        want = s1 - 0.05 * math.sqrt(1 - 0.999 ** 2) / (1 - 0.9 ** 2) * m2 / (v2.sqrt() + 1e-8)
    assert torch.allclose(params["w"], want, atol=1e-6)
    assert opt.global_step == 2
    # EMA with num_updates: d0 = 0.1, d1 = 2/11
    e1 = w0 - 0.9 * (w0 - (want if False else None)) if False else None  # noqa: F841 (formula spelled out below)
    first = {"sgd": w0 - 0.1 * g, "momentum": w0 - 0.1 * g}.get(kind)
    if first is not None:
        ema1 = w0 - (1 - 0.1) * (w0 - first)
        ema2 = ema1 - (1 - 2 / 11) * (ema1 - want)
        assert torch.allclose(opt.ema["w"], ema2, atol=1e-6)


# ---- engine + checkpoint ---------------------------------------------------------------------------------------
def _torch_engine(preset="reference", B=8, **kw):
    from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

    spec = presets.make_spec(preset)
    eng = TorchEngine(spec, B, optimizer=kw.pop("optimizer", "adam"), base_lr=kw.pop("base_lr", 1e-3), **kw)
    eng.load_params(oracle.init_params(spec, seed=0))
    return spec, eng


def test_philox_mirror_known_answers_and_mask_statistics():
    """The numpy mirror of the dropout generator (csrc/common/philox.cuh) against the Random123 known-answer vectors of
    philox4x32-10; the GPU tests then require the kernels' masks to equal this mirror bit for bit."""
    from distributed_tensorflow_ibm_mnist_b200.ops.philox import dropout_keep_mask, philox4x32_10

    u = lambda v: np.array([v], dtype=np.uint32)
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = philox4x32_10(*[u(c) for c in ctr], np.uint32(key[0]), np.uint32(key[1]))
        assert tuple(int(g[0]) for g in got) == want
    m1 = dropout_keep_mask(64, 256, seed=7, step=3, keep_prob=0.5)
    assert m1.shape == (64, 256) and m1.dtype == np.bool_ and 0.45 < m1.mean() < 0.55
    assert np.array_equal(m1, dropout_keep_mask(64, 256, seed=7, step=3, keep_prob=0.5))          # deterministic
    assert not np.array_equal(m1, dropout_keep_mask(64, 256, seed=7, step=4, keep_prob=0.5))      # step dependent
    assert dropout_keep_mask(8, 64, seed=1, step=0, keep_prob=0.9).mean() > 0.8


def test_synthetic_batches_are_packed_for_single_copy_staging():
    """Pixels and labels of a batch share one buffer (pixels | pad to 16 B | int32 labels): the feeder's one-copy path."""
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import DeviceFeeder, SyntheticMnist

    for B, px in ((6, 784), (5, 2352), (3, 787)):
        src = SyntheticMnist(B, px, pool_batches=3, seed=4, pin=False)
        nb = B * px
        assert src.label_offset % 16 == 0 and nb <= src.label_offset < nb + 16
        assert src.packed.shape == (3, src.label_offset + 4 * B)
        x, y = src.next()
        buf = src.packed[0]
        assert x.shape == (B, px) and y.shape == (B,) and y.dtype == torch.int32
        assert torch.equal(buf[:nb].view(B, px), x) and torch.equal(buf[src.label_offset:].view(torch.int32), y)
        assert torch.equal(src.next_packed(), src.packed[1])          # the cursor is shared by both accessors

    class _Eng:       # the feeder only takes the packed path when the engine's slot layout matches the source's
        B, device = 6, torch.device("cpu")
        packed_label_offset = 6 * 784
        class spec:
            image_pixels = 784
        def stage_packed(self, *a, **k): ...
    if torch.cuda.is_available():
        assert DeviceFeeder(_Eng(), SyntheticMnist(6, 784, pool_batches=2, pin=False))._packed


def test_torch_engine_learns_separable_data():
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import SyntheticMnist

    spec, eng = _torch_engine("baseline", B=32, optimizer="momentum", base_lr=0.05)
    src = SyntheticMnist(32, spec.image_pixels, pool_batches=4, seed=0, separable=True, pin=False)
    first = last = None
    for i in range(30):
        eng.stage_batch(0, *src.next())
        eng.train_step(0)
        m = eng.metrics()
        first = first if first is not None else m["cross_entropy"]
        last = m["cross_entropy"]
    assert last < 0.5 * first and m["global_step"] == 30 and not m["nan"]
    probs = eng.predict(src.images[0])
    assert probs.shape == (32, 10) and torch.allclose(probs.sum(1), torch.ones(32), atol=1e-5)


def test_checkpoint_inventory_resume_and_rotation(tmp_path):
    from distributed_tensorflow_ibm_mnist_b200.formats import checkpoint_state as cs
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing as ck

    spec, eng = _torch_engine("reference", B=4)
    x = torch.randint(0, 256, (4, spec.image_pixels), dtype=torch.uint8)
    y = torch.randint(0, 10, (4,), dtype=torch.int32)
    d = str(tmp_path / "train")
    prefixes = []
    for i in range(7):
        eng.stage_batch(0, x, y)
        eng.train_step(0)
        prefixes.append(ck.save_checkpoint(d, eng, spec, "adam", num_shards=2, max_to_keep=5))
    rd = BundleReader(prefixes[-1])
    keys = set(rd.keys())
    expect = {"global_step", "beta1_power", "beta2_power", "cross_entropy/avg", "total_loss/avg"}
    for v in spec.variables:
        expect |= {v.name, v.name + "/ExponentialMovingAverage", v.name + "/Adam", v.name + "/Adam_1"}
    expect |= {f"{l}/weight_loss/avg" for l in ("conv1", "conv2", "local3", "local4", "softmax_linear")}
    assert keys == expect
    assert rd.num_shards == 2 and int(rd.get("global_step")) == 7 and rd.entry("global_step").dtype == 9
    assert rd.entry("conv1/weights/Adam").shard_id == rd.entry("conv1/weights").shard_id      # slots co-located
    st = cs.read_state(d)
    assert st.model_checkpoint_path == "model.ckpt-7" and st.all_model_checkpoint_paths == [f"model.ckpt-{i}" for i in range(3, 8)]
    assert not os.path.exists(prefixes[0] + ".index") and os.path.exists(prefixes[-1] + ".index")
    # every checkpoint carries a MetaGraphDef (Saver(write_meta_graph=True)); rotation removes it with the bundle
    assert os.path.exists(prefixes[-1] + ".meta") and not os.path.exists(prefixes[0] + ".meta")
    from tensorboard.compat.proto import meta_graph_pb2
    mg = meta_graph_pb2.MetaGraphDef()
    mg.ParseFromString(open(prefixes[-1] + ".meta", "rb").read())
    assert {n.name for n in mg.graph_def.node} == expect and all(n.op == "VariableV2" for n in mg.graph_def.node)
    assert mg.saver_def.version == 2 and mg.saver_def.sharded and mg.saver_def.max_to_keep == 5
    assert set(ck.variable_inventory(spec, "adam")) == expect        # what graph.pbtxt lists, derived without an engine
    # resume into a fresh engine: bit-exact variables, optimizer state and step
    spec2, eng2 = _torch_engine("reference", B=4)
    assert ck.restore_checkpoint(d, eng2, spec2, "adam") == 7
    for k, v in eng.params().items():
        assert torch.equal(v, eng2.params()[k])
    for a, b in zip(eng.slots(), eng2.slots()):
        for k in a:
            assert torch.equal(a[k], b[k])
    for k, v in eng.ema_params().items():
        assert torch.equal(v, eng2.ema_params()[k])
    assert eng2.metrics()["global_step"] == 7
    eng.stage_batch(0, x, y); eng.train_step(0)
    eng2.stage_batch(0, x, y); eng2.train_step(0)
    for k, v in eng.params().items():
        assert torch.allclose(v, eng2.params()[k], atol=1e-7), k
    assert ck.restore_checkpoint(str(tmp_path / "nothing"), eng2, spec2, "adam") is None


def test_checkpoint_channel_adaptation_for_inference(tmp_path):
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing as ck

    spec3, eng3 = _torch_engine("reference", B=4)           # 3-channel training graph
    d = str(tmp_path)
    ck.save_checkpoint(d, eng3, spec3, "adam")
    spec1 = presets.make_spec("reference", channels=1)
    from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

    eng1 = TorchEngine(spec1, 4)
    ck.restore_checkpoint(d, eng1, spec1, "sgd", model_only=True)
    grey = torch.randint(0, 256, (4, 784), dtype=torch.uint8)
    rgb = grey.view(4, 784, 1).repeat(1, 1, 3).view(4, 2352)
    assert torch.allclose(eng1.predict(grey), eng3.predict(rgb), atol=1e-5)


# ---- data ------------------------------------------------------------------------------------------------
def _write_idx(tmp, n=30):
    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, (n, 28, 28), dtype=np.uint8)
    labs = rng.integers(0, 10, n).astype(np.uint8)
    for split, (fi, fl) in {"training": ("train-images.idx3-ubyte", "train-labels.idx1-ubyte"),
                            "testing": ("t10k-images.idx3-ubyte", "t10k-labels.idx1-ubyte")}.items():
        with open(os.path.join(tmp, fi), "wb") as f:
            f.write(struct.pack(">IIII", 2051, n, 28, 28) + imgs.tobytes())
        with open(os.path.join(tmp, fl), "wb") as f:
            f.write(struct.pack(">II", 2049, n) + labs.tobytes())
    return imgs, labs


def test_graph_pbtxt_and_step_counter(tmp_path):
    """The chief's session-start artefacts: graph.pbtxt parses as a GraphDef with the variable inventory, and
    the step counter writes global_step/sec (+ images/sec) into the train_dir event file every N steps."""
    from google.protobuf import text_format
    from tensorboard.compat.proto import graph_pb2

    from distributed_tensorflow_ibm_mnist_b200.formats import events, graph_def as gd
    from distributed_tensorflow_ibm_mnist_b200.models import presets
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing as ck, trainer

    spec = presets.make_spec("baseline")
    d = str(tmp_path / "train")
    path = ck.write_graph(d, spec, "momentum")
    g = graph_pb2.GraphDef()
    text_format.Parse(open(path).read(), g)
    inv = ck.variable_inventory(spec, "momentum")
    assert [n.name for n in g.node] == list(inv)
    node = {n.name: n for n in g.node}
    assert [dim.size for dim in node["local3/weights/Momentum"].attr["shape"].shape.dim] == [3136, 1024]
    assert node["global_step"].attr["dtype"].type == 9 and node["conv1/biases"].attr["dtype"].type == 1
    g2 = graph_pb2.GraphDef()
    g2.ParseFromString(gd.graph_def(inv))                       # binary encoder agrees with the text one
    assert g2 == g

    class _Loop:
        global_step = 0

    w = events.EventFileWriter(d)
    hook = trainer.StepCounterHook(w, every_n_steps=5, global_batch=64)
    loop = _Loop()
    hook.begin(loop)
    ctx = trainer.RunContext(loop)
    for _ in range(12):
        loop.global_step += 1
        hook.after_run(ctx)
    hook.end(loop)
    w.close()
    evs = [e for f in sorted(os.listdir(d)) if "tfevents" in f for e in events.read_events(os.path.join(d, f))]
    tags = [(e["step"], tag) for e in evs for tag in e["scalars"]]
    assert (5, "global_step/sec") in tags and (10, "global_step/sec") in tags and (10, "images/sec") in tags
    assert not any(step == 12 for step, _ in tags) and hook.last_rate > 0


def test_collective_saves_over_gloo(tmp_path):
    """Sharded engines need every rank inside ``save`` (the optimizer state is gathered from its owners): the chief's
    time-based decision is broadcast at fixed step intervals and all ranks act on it together; only the chief writes."""
    import json
    import socket
    import subprocess
    import sys

    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    worker = os.path.join(os.path.dirname(__file__), "helpers", "collective_save_worker.py")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, worker, str(r), "2", str(port), str(tmp_path)], env=env,
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240) for p in procs]
    assert all(p.returncode == 0 for p in procs), [o[1][-1500:] for o in outs]
    c0 = json.load(open(tmp_path / "calls0.json"))["calls"]
    c1 = json.load(open(tmp_path / "calls1.json"))["calls"]
    assert [c[0] for c in c0] == [c[0] for c in c1] == [4, 8, 10]          # two polled saves + the final one, in lock step
    assert all(w for _, w in c0) and not any(w for _, w in c1)              # only the chief writes
    files = sorted(os.listdir(tmp_path / "train"))
    assert "model.ckpt-10.index" in files and "model.ckpt-4.index" in files and "checkpoint" in files


def test_convert_mnist_idx_to_png_to_tfrecord_roundtrip(tmp_path):
    sys.path.insert(0, ROOT)
    import convert_mnist
    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord

    src = tmp_path / "idx"; src.mkdir()
    imgs, labs = _write_idx(str(src))
    lbl, data, size, rows, cols = convert_mnist.read("training", str(src))
    assert (size, rows, cols) == (30, 28, 28) and list(lbl) == list(labs)
    with pytest.raises(ValueError):
        convert_mnist.read("validation", str(src))
    out = tmp_path / "png"
    assert convert_mnist.main(["convert_mnist.py", str(src), str(out), "--quiet"]) == 0
    f0 = out / "training" / str(labs[0]) / "0.png"
    assert f0.exists() and np.array_equal(readers.load_image_gray(str(f0)), imgs[0])
    rec = tmp_path / "train.tfrecord"
    assert convert_mnist.main(["convert_mnist.py", "png2tfrecord", str(out / "training"), str(rec), "--channels=3"]) == 0
    exs = [tfrecord.decode_example(r) for r in tfrecord.read_records(str(rec))]
    assert len(exs) == 30 and all(len(e["image_raw"]) == 2352 for e in exs)
    by_label = sorted(labs.tolist())
    assert [e["label"] for e in exs] == by_label
    first = np.frombuffer(exs[0]["image_raw"], np.uint8).reshape(28, 28, 3)
    assert np.array_equal(first[..., 0], first[..., 2])
    rec2 = tmp_path / "train_space.tfrecord"      # the space-separated spelling of the same option
    assert convert_mnist.main(["convert_mnist.py", "png2tfrecord", str(out / "training"), "--channels", "3", str(rec2)]) == 0
    assert rec2.read_bytes() == rec.read_bytes()
    assert convert_mnist.main(["convert_mnist.py", "png2tfrecord", str(out / "training"), str(rec2), "--channels", "2"]) == 2
    assert convert_mnist.main(["convert_mnist.py"]) == 2


@pytest.mark.parametrize("native", [False, True])
def test_tfrecord_batches_shuffle_and_validation(tmp_path, native):
    """Both implementations of the training input pipeline: Python threads and the C++ loader (record_loader.cpp)."""
    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord

    if native and readers._native_loader_lib() is None:
        pytest.skip("libdtm_b200.so not built")

    p = str(tmp_path / "d.tfrecord")
    n = 300
    imgs = [bytes([i % 256]) * 784 for i in range(n)]
    tfrecord.write_mnist_tfrecord(p, imgs, [i % 10 for i in range(n)])
    with pytest.raises(ValueError, match="Failed to find file"):
        readers.TFRecordBatches([p, str(tmp_path / "missing")], 8, 784, native=native)
    ds = readers.TFRecordBatches([p], 16, 784, min_after_dequeue=50, seed=1, native=native)
    assert isinstance(ds, readers.NativeRecordBatches) == native
    assert ds.capacity == 50 + 3 * 16
    x, y = ds.next()
    assert x.shape == (16, 784) and x.dtype == torch.uint8 and y.dtype == torch.int32
    assert all(int(x[i, 0]) % 10 == int(y[i]) for i in range(16))           # image/label pairs stay together
    seen = [int(v) for _ in range(10) for v in ds.next()[0][:, 0]]
    assert seen != sorted(seen)                                              # shuffled
    ds.close()
    bad = readers.TFRecordBatches([p], 4, 2352, min_after_dequeue=0, native=native)
    with pytest.raises(ValueError, match="expected 2352"):
        bad.next()
    bad.close()
    # two reader threads share ONE record stream (the reference's single TFRecordReader): nothing of the first epoch is
    # delivered twice.  1200 distinct records (id in the first two pixels); the pipeline looks ahead at most one chunk of 128
    # records per thread + capacity + min_after + batch = 256 + 50 + 20 + 10 records, so the first 800 dequeued records all
    # belong to epoch one
    pu = str(tmp_path / "unique.tfrecord")
    tfrecord.write_mnist_tfrecord(pu, [bytes([i % 256, i // 256]) + bytes(782) for i in range(1200)], [i % 10 for i in range(1200)])
    once = readers.TFRecordBatches([pu], 10, 784, min_after_dequeue=20, seed=4, num_threads=2, native=native)
    got = []
    for _ in range(80):
        xb = once.next()[0]
        got += [int(a) + 256 * int(b) for a, b in zip(xb[:, 0], xb[:, 1])]
    assert len(set(got)) == len(got) == 800
    once.close()
    # unshuffled: file order, every record exactly once per epoch
    seq = readers.TFRecordBatches([p], 50, 784, shuffle=False, native=native)
    first = [int(v) for _ in range(6) for v in seq.next()[0][:, 0]]
    assert first == [i % 256 for i in range(300)]
    seq.close()
    if native:
        # a flipped payload byte must be caught by the record CRC, and the packed layout must match the engine's slots
        raw = bytearray(open(p, "rb").read())
        raw[40] ^= 0xFF
        q = str(tmp_path / "corrupt.tfrecord")
        open(q, "wb").write(bytes(raw))
        c = readers.TFRecordBatches([q], 4, 784, min_after_dequeue=0, native=True)
        with pytest.raises(ValueError, match="corrupt TFRecord"):
            c.next()
        c.close()
        ds2 = readers.TFRecordBatches([p], 16, 784, min_after_dequeue=10, seed=3, native=True)
        buf = ds2.next_packed()
        assert buf.numel() == ds2.packed_bytes == 16 * 784 + 4 * 16 and ds2.label_offset == 16 * 784
        lab = buf[ds2.label_offset:].view(torch.int32)
        assert all(int(buf[i * 784]) % 10 == int(lab[i]) for i in range(16))
        ds2.close()
    got = list(readers.iter_tfrecord_examples([p]))
    assert len(got) == n and got[3][2] == 3 and got[0][0].endswith(":0")


def test_native_loader_mixed_record_shapes(tmp_path):
    """The C++ reader's remembered-record-shape fast path must fall back to the general Example parser whenever a record
    differs from the remembered one anywhere outside its pixels and label value: other feature sets, other field order,
    labels that need a two-byte varint, un-packed int64 lists.  Every record is checked against the Python codec."""
    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.formats import proto, tfrecord

    if readers._native_loader_lib() is None:
        pytest.skip("libdtm_b200.so not built")

    def bytes_feature(b):
        return proto.f_bytes(1, proto.f_bytes(1, b))

    def int_feature(v, packed=True):
        return proto.f_bytes(3, proto.f_bytes(1, proto.enc_varint(v)) if packed else proto.f_varint(1, v))

    def example(entries):
        return proto.f_bytes(1, b"".join(proto.f_bytes(1, proto.f_str(1, k) + proto.f_bytes(2, f)) for k, f in entries))

    rng = np.random.default_rng(5)
    recs, want = [], []
    for i in range(400):
        img = rng.integers(0, 256, 784, dtype=np.uint8).tobytes()
        label = [i % 10, 128 + i % 100, 5, 300][i % 4] if i % 7 == 0 else i % 10
        kind = (i // 5) % 4                                  # runs of five records per structure
        if kind == 0:
            e = example([("image_raw", bytes_feature(img)), ("label", int_feature(label))])
        elif kind == 1:
            e = example([("label", int_feature(label)), ("image_raw", bytes_feature(img))])
        elif kind == 2:
            e = example([("depth", int_feature(1)), ("image_raw", bytes_feature(img)), ("label", int_feature(label, packed=False)),
                         ("width", int_feature(28))])
        else:
            e = tfrecord.mnist_example(img, label)
        assert tfrecord.decode_example(e)["label"] == label
        recs.append(e)
        want.append((img, label))
    path = str(tmp_path / "mixed.tfrecord")
    with tfrecord.TFRecordWriter(path) as w:
        for e in recs:
            w.write(e)
    ds = readers.TFRecordBatches([path], 50, 784, shuffle=False, native=True)
    for b in range(8):
        x, y = ds.next()
        for j in range(50):
            img, label = want[b * 50 + j]
            assert bytes(x[j].numpy().tobytes()) == img and int(y[j]) == label, (b, j)
    ds.close()


def test_native_loader_batch_ring_lends_buffers(tmp_path):
    """Batch ring of the C++ loader: batches arrive in the same order as from the caller-thread dequeue, a lent buffer is
    not rewritten during the next prefetch_batches - 2 calls, and the ring and the direct call exclude each other."""
    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord

    lib = readers._native_loader_lib()
    if lib is None:
        pytest.skip("libdtm_b200.so not built")
    p = str(tmp_path / "d.tfrecord")
    tfrecord.write_mnist_tfrecord(p, [bytes([i % 256, i // 256]) + bytes(782) for i in range(700)], [i % 10 for i in range(700)])

    def ids(buf, ds):
        px = buf[:ds.B * 784].view(ds.B, 784)
        return [int(a) + 256 * int(b) for a, b in zip(px[:, 0], px[:, 1])]

    ring = readers.TFRecordBatches([p], 20, 784, shuffle=False, native=True, prefetch_batches=4)
    direct = readers.TFRecordBatches([p], 20, 784, shuffle=False, native=True, prefetch_batches=0)
    held = []
    for b in range(40):                                  # 800 records: runs into the second epoch
        buf = ring.next_packed()
        want = [(b * 20 + j) % 700 for j in range(20)]
        assert ids(buf, ring) == want == ids(direct.next_packed(), direct)
        assert [int(v) for v in buf[ring.label_offset:].view(torch.int32)] == [w % 10 for w in want]
        held.append((buf, want))
        for old, old_want in held[-3:]:                  # this buffer and the two lent before it are still intact
            assert ids(old, ring) == old_want
    assert len({h[0].data_ptr() for h in held}) == 4     # four buffers go round
    assert lib.dtm_loader_next(ring._h, held[0][0].data_ptr(), held[0][0].data_ptr() + ring.label_offset) != 0
    assert b"ring" in lib.dtm_last_error()
    ring.close()
    direct.close()
    # shuffled through the ring: pairs stay together, nothing of the first epoch twice (the pipeline looks ahead at most
    # 2 x 128 chunk records + capacity 88 + ring 48 + batch 16 = 408 records: 240 delivered + 408 < 700)
    sh = readers.TFRecordBatches([p], 16, 784, seed=9, min_after_dequeue=40, num_threads=2, native=True, prefetch_batches=3)
    got = []
    for _ in range(15):
        buf = sh.next_packed()
        i = ids(buf, sh)
        assert [int(v) for v in buf[sh.label_offset:].view(torch.int32)] == [k % 10 for k in i]
        got += i
    assert len(set(got)) == len(got) == 240 and got != sorted(got)
    sh.close()


def test_inference_decode_stage_keeps_order():
    """inference.py's optional decode threads: results come back in input order whatever the completion order."""
    import time

    import inference

    def slow_square(i):
        time.sleep(0.0005 * (i % 4))
        return i * i

    for threads in (1, 4):
        assert list(inference._ordered_parallel_map(slow_square, iter(range(150)), threads, depth=16)) == [i * i for i in range(150)]
    with pytest.raises(ZeroDivisionError):
        list(inference._ordered_parallel_map(lambda i: 1 // (i - 5), range(10), 3))


def test_image_preprocessing_helpers():
    from distributed_tensorflow_ibm_mnist_b200.data import readers

    img = np.arange(30 * 20, dtype=np.uint8).reshape(30, 20)
    out = readers.crop_or_pad(img, 28)
    assert out.shape == (28, 28) and np.array_equal(out[:, 4:24], img[1:29]) and out[:, :4].sum() == 0
    s = readers.per_image_standardization(np.random.default_rng(0).integers(0, 255, (28, 28)).astype(np.uint8))
    assert abs(s.mean()) < 1e-5 and abs(s.std() - 1) < 1e-4
    flat = readers.per_image_standardization(np.full((28, 28), 7, np.uint8))
    assert np.all(flat == 0)                                                  # stddev floor 1/sqrt(N)


def test_monitor_collects_reference_summaries():
    from distributed_tensorflow_ibm_mnist_b200.formats import proto
    from distributed_tensorflow_ibm_mnist_b200.train import monitor as mc

    mon = mc.CMonitor("/tmp/x", 10, 100)
    for layer in ["conv1", "local3"]:
        mon.SummaryHist("weight", layer + "/weights:0", layer)
        mon.SummaryHist("activation", layer + "/" + layer + ":0", layer)
        mon.SummaryNorm2("weight", layer + "/weights:0", layer)
    mon.SummaryGradient("weight", "loss")
    mon.SummaryGWRatio()
    mon.SummaryScalar("train loss", "loss"); mon.SummaryScalar("test accuracy", "accuracy")
    t = {"params": {"conv1/weights": torch.ones(5, 5, 1, 32), "local3/weights": torch.full((4, 4), 2.0)},
         "activations": {"conv1": torch.zeros(2, 3), "local3": torch.ones(2, 3)},
         "gradients": {"conv1/weights": torch.full((5, 5, 1, 32), 0.5), "local3/weights": torch.ones(4, 4)}}
    tags = {}
    for f, _, v in proto.parse(mon.collect(mc.DLMAO_TRAIN_SUMMARIES, t, {"loss": 1.5, "accuracy": 0.25})):
        d = proto.parse_dict(v)
        tags[d[1][0].decode()] = d
    assert {"train loss", "conv1/weight", "conv1/activation", "conv1/weight_norm2", "local3/weight_norm2",
            "conv1/weight_gradient", "conv1/gradient_weight_ratio", "local3/gradient_weight_ratio"} <= set(tags)
    norm = struct.unpack("<f", struct.pack("<I", tags["local3/weight_norm2"][2][0]))[0]
    ratio = struct.unpack("<f", struct.pack("<I", tags["conv1/gradient_weight_ratio"][2][0]))[0]
    assert abs(norm - 8.0) < 1e-5 and abs(ratio - 0.5) < 1e-5
    test_tags = [proto.parse_dict(v)[1][0].decode() for f, _, v in proto.parse(mon.collect(mc.DLMAO_TEST_SUMMARIES, t, {"loss": 1.0, "accuracy": 0.5}))]
    assert test_tags == ["test accuracy"]


def test_inference_helper_result_files(tmp_path):
    from distributed_tensorflow_ibm_mnist_b200.utils import inference_helper as ih

    probs = np.array([[0.7, 0.2, 0.1] + [0] * 7, [0.05, 0.9, 0.05] + [0] * 7], np.float32)
    lab = tmp_path / "labels.txt"
    lab.write_text("\n".join(f"digit{i}" for i in range(10)))
    doc = ih.writeClassificationResult(str(tmp_path / "o" / "r.json"), ["a.png", "b.png"], probs, prob_thresh=0.15, label_file=str(lab))
    assert [p["label"] for p in doc["results"][0]["predictions"]] == ["digit0", "digit1"]
    assert json.load(open(tmp_path / "o" / "r.json"))["num_samples"] == 2
    doc = ih.writeClassificationResult(str(tmp_path / "v.json"), ["k0", "k1"], probs, ground_truth=[0, 3])
    assert doc["accuracy"] == 0.5 and doc["results"][1]["correct"] is False


def test_thread_plane_collectives():
    """Control plane of the single-process tower mode: threads instead of ranks (parallel/control_plane.py)."""
    import threading

    from distributed_tensorflow_ibm_mnist_b200.parallel.control_plane import ThreadPlane, as_plane

    planes = ThreadPlane.make(3)
    out = [None] * 3

    def run(r):
        p = planes[r]
        assert as_plane(p) is p and p.same_process and p.world_size == 3
        a = p.all_gather_object({"rank": r})
        b = p.broadcast_object("from-2" if r == 2 else None, src=2)
        t = torch.full((4,), float(r))
        p.broadcast_tensor(t, 1)
        p.barrier()
        out[r] = (a, b, t.tolist(), p.all_true(r >= 0), p.all_true(r > 0), p.any_true(r == 2), p.any_true(False))

    ths = [threading.Thread(target=run, args=(r,)) for r in range(3)]
    for t in ths:
        t.start()
    for t in ths:
        t.join(timeout=60)
    for r in range(3):
        a, b, t, all1, all2, any1, any2 = out[r]
        assert a == [{"rank": 0}, {"rank": 1}, {"rank": 2}] and b == "from-2" and t == [1.0] * 4
        assert all1 and not all2 and any1 and not any2


def test_nan_hook_polls_sparsely_and_stops_all_ranks_together():
    """NanLossHook (the reference's NanTensorHook(loss), main.py:138): the engines keep a sticky NaN flag, the hook polls it
    every ``check_every`` steps and at the end; with several ranks the decision is collective - a rank whose own loss is
    still finite stops on the same step as the rank that saw the NaN, instead of waiting for it in the next exchange."""
    import threading

    from distributed_tensorflow_ibm_mnist_b200.parallel.control_plane import ThreadPlane
    from distributed_tensorflow_ibm_mnist_b200.train import trainer

    class Engine:                                   # NaN appears at local step `bad_at` and sticks
        def __init__(self, bad_at):
            self.bad_at, self.steps, self.polls = bad_at, 0, 0

        def stage_batch(self, slot, x, y):
            pass

        def train_step(self, slot):
            self.steps += 1

        def metrics(self):
            self.polls += 1
            bad = self.bad_at is not None and self.steps >= self.bad_at
            return {"global_step": self.steps, "total_loss": float("nan") if bad else 1.0, "nan": bad}

    class Source:
        def next(self):
            return None, None

    def run(engine, hook, plane=None, steps=1000):
        with trainer.MonitoredTrainingLoop(engine, Source(), hooks=[trainer.StopAtStepHook(steps), hook], group=plane,
                                           log_every=0, out=lambda *a: None) as sess:
            while not sess.should_stop():
                sess.run_step()

    assert trainer.NanLossHook().check_every == 100
    eng = Engine(bad_at=130)
    with pytest.raises(trainer.NanLossError, match="NaN loss"):
        run(eng, trainer.NanLossHook())
    assert eng.steps == 200 and eng.polls == 1 + 2           # start-step read, polls at 100 and 200; stops at the second
    eng = Engine(bad_at=7)
    run(eng, trainer.NanLossHook(check_every=5, fail_on_nan_loss=False))
    assert eng.steps == 10                                    # request_stop instead of raising
    eng = Engine(bad_at=None)
    run(eng, trainer.NanLossHook(), steps=250)
    assert eng.steps == 250 and eng.polls == 1 + 2 + 1        # ... and the final check in end()

    planes = ThreadPlane.make(2)
    engines = [Engine(bad_at=None), Engine(bad_at=30)]
    errors = [None, None]

    def worker(r):
        try:
            run(engines[r], trainer.NanLossHook(check_every=20), plane=planes[r], steps=500)
        except trainer.NanLossError as e:
            errors[r] = e

    ths = [threading.Thread(target=worker, args=(r,)) for r in range(2)]
    for t in ths:
        t.start()
    for t in ths:
        t.join(timeout=60)
    assert all(isinstance(e, trainer.NanLossError) for e in errors)
    assert engines[0].steps == engines[1].steps == 40         # both at the first poll after the event


def test_dist_plane_passes_file_descriptors(tmp_path):
    """How GPU allocation handles travel between worker processes: SCM_RIGHTS over abstract unix sockets, rendezvous over the
    process group (gloo here)."""
    import socket as _s
    import subprocess

    with _s.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    worker = os.path.join(ROOT, "tests", "helpers", "fd_exchange_worker.py")
    procs = [subprocess.Popen([sys.executable, worker, str(r), "2", str(port), str(tmp_path)], stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True, env=dict(os.environ, PYTHONPATH=ROOT)) for r in range(2)]
    outs = [p.communicate(timeout=180) for p in procs]
    for r, (p, (o, e)) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"FD_EXCHANGE_OK {r}" in o, e[-2000:]


def test_kernel_register_budgets_from_the_build_log():
    """Performance invariants that live in register counts (ptxas -v output kept by the build): kernels that are designed to
    SHARE an SM must keep fitting together, the LRN kernels must stay spill free, and no new spills may appear unnoticed."""
    import re

    log = os.path.join(ROOT, "distributed_tensorflow_ibm_mnist_b200", "_native", "ptxas.log")
    if not os.path.exists(log):
        pytest.skip("library not built here")
    text = open(log).read()
    kernels = {}
    for m in re.finditer(r"Compiling entry function '(\w+)' for 'sm_100a'.*?(\d+) bytes spill stores, (\d+) bytes spill loads.*?Used (\d+) registers",
                         text, flags=re.S):
        kernels[m.group(1)] = {"spill": int(m.group(2)) + int(m.group(3)), "regs": int(m.group(4))}
    assert len(kernels) >= 40, len(kernels)

    def find(*parts):
        hits = [k for k in kernels if all(p in k for p in parts)]
        assert len(hits) == 1, (parts, hits)
        return kernels[hits[0]]

    def alloc(regs, threads):                  # registers are allocated per warp in units of 8 per thread
        return -(-regs // 8) * 8 * threads

    update = find("fused_update_stride_kernel")
    dgrad, dgrad_lrn = find("conv2_dgrad_kernelILb0"), find("conv2_dgrad_kernelILb1")
    # the early optimizer update (384 threads) runs beside conv2_dgrad (384 threads) on every SM: 64 K registers per SM
    assert alloc(update["regs"], 384) + alloc(dgrad["regs"], 384) <= 65536
    assert alloc(update["regs"], 384) + alloc(dgrad_lrn["regs"], 384) <= 65536
    # two 320-thread conv1 CTAs per SM (one image channel)
    assert 2 * alloc(find("conv1_sw_kernelILi1ELb0")["regs"], 320) <= 65536
    # spill free: everything except the conv1 kernels' known 8-24 byte frames
    for name, k in kernels.items():
        if "conv1_tc_kernel" in name or "conv1_sw_kernelILi1" in name:
            assert k["spill"] <= 80, (name, k)
        else:
            assert k["spill"] == 0, (name, k)
    assert find("unpool2_lrn_kernel")["regs"] <= 64


def test_built_library_is_blackwell_native_sass():
    """SASS of the built library (cuobjdump, no GPU needed): every GEMM-shaped kernel issues tcgen05.mma (UTC*MMA) with TMEM
    read-back (LDTM) and bulk/TMA operand loads, nothing falls back to legacy mma.sync (HMMA), and the fused exchange
    kernels carry multimem (NVLS) reductions."""
    import re
    import shutil
    import subprocess

    lib = os.path.join(ROOT, "distributed_tensorflow_ibm_mnist_b200", "_native", "libdtm_b200.so")
    if not os.path.exists(lib) or shutil.which("cuobjdump") is None:
        pytest.skip("library or cuobjdump not available")
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, timeout=300).stdout
    parts = re.split(r"^\s*Function : (\S+)\s*$", txt, flags=re.M)
    funcs = {parts[k]: parts[k + 1] for k in range(1, len(parts), 2)}
    assert len(funcs) >= 40

    def count(body, mnemonic):
        return len(re.findall(r"\b" + mnemonic, body))

    assert all(count(b, "HMMA") == 0 for b in funcs.values())
    tensor = [n for n in funcs if any(t in n for t in ("gemm_bf16_kernel", "conv2_fwd_kernel", "conv2_dgrad_kernel",
                                                       "conv_wgrad_fused_kernel", "conv1_sw_kernel", "conv2_wgrad_kernel"))]
    assert len(tensor) >= 15
    for n in tensor:
        b = funcs[n]
        assert count(b, "UTC[A-Z]*MMA") > 0 and count(b, "LDTM") > 0, n
        assert count(b, "UBLKCP") + count(b, "UTMALDG") + count(b, "STS") > 0, n
    assert all(count(funcs[n], "UTMALDG") > 0 for n in funcs if "gemm_bf16_kernel" in n)          # TMA tensor maps
    assert all(count(funcs[n], "UBLKCP") > 0 for n in funcs if "conv2_fwd_kernel" in n or "conv_wgrad_fused" in n)
    nvls = [n for n in funcs if "fused_allreduce_update_kernelILb1" in n or "allreduce_kernel" in n]
    assert nvls and all(count(funcs[n], "LDGMC") > 0 for n in nvls if "fused_allreduce_update_kernelILb1" in n)


def test_no_cluster_multicast_loads_in_peer_memory_kernels():
    """TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung
    the GPU: the communication kernels (the only code that holds peer pointers) contain no cluster-multicast bulk load."""
    import re

    comm = os.path.join(ROOT, "csrc", "comm")
    for name in sorted(os.listdir(comm)):
        text = open(os.path.join(comm, name)).read()
        assert not re.search(r"cp\.async\.bulk[\w.:]*\.multicast::cluster", text), name


def test_c_abi_header_library_and_bindings_agree():
    """include/dtm_b200.h is the contract: every function it declares is exported by the built library, and every ctypes
    signature the Python bindings register has the header's parameter count (the mismatch ctypes cannot detect itself)."""
    import ctypes
    import re

    from distributed_tensorflow_ibm_mnist_b200.ops import _lib

    if not _lib.available():
        pytest.skip("library not built here")
    text = open(os.path.join(ROOT, "include", "dtm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    decls = {}
    for m in re.finditer(r"\b(dtm_\w+)\s*\(([^()]*)\)\s*;", text):
        args = m.group(2).strip()
        decls[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    assert len(decls) >= 50 and decls["dtm_abi_version"] == 0 and decls["dtm_loader_ring_start"] == 4
    assert int(re.search(r"#define DTM_ABI_VERSION (\d+)", open(os.path.join(ROOT, "include", "dtm_b200.h")).read()).group(1)) == _lib.ABI_VERSION
    lib = ctypes.CDLL(str(_lib.lib_path()))
    missing = [n for n in decls if not hasattr(lib, n)]
    assert not missing, missing
    # importing the op modules registers their signatures
    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.ops import native  # noqa: F401
    from distributed_tensorflow_ibm_mnist_b200.parallel import peer_comm  # noqa: F401

    readers._native_loader_lib()
    assert len(_lib.EXTRA_SIGNATURES) >= 30
    for name, (argtypes, _res) in _lib.EXTRA_SIGNATURES.items():
        assert name in decls, f"{name} is bound from Python but not declared in dtm_b200.h"
        assert len(argtypes) == decls[name], (name, len(argtypes), decls[name])
    loaded = _lib.load()                     # ... and whatever else the loader itself declares (gemm, probe, device info)
    bound = 0
    for name, n_args in decls.items():
        fn = getattr(loaded, name)
        if fn.argtypes is not None:
            bound += 1
            assert len(fn.argtypes) == n_args, (name, len(fn.argtypes), n_args)
    assert bound >= 35


def test_ctypes_struct_mirrors_match_the_c_header(tmp_path):
    """dtm_train_state / dtm_update_config cross the ABI by value and by device copy: the ctypes mirrors in ops/native.py must
    have the C compiler's size and field offsets (a plain-C translation unit including the header also proves the header is C)."""
    import ctypes
    import shutil
    import subprocess

    from distributed_tensorflow_ibm_mnist_b200.ops import native

    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    fields_state = ["global_step", "nan_flag", "ticket", "last_xent", "wl_sum", "last_wl", "loss_avg", "work"]
    fields_cfg = ["optimizer", "decay_steps", "base_lr", "ema_decay", "seg_offset", "seg_wd", "seg_zero_grad", "seg_gather_bf16",
                  "pack_seg", "pack_fwd", "pack_dgrad"]
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "dtm_b200.h"\nint main(void) {\n'
                   '  printf("%zu %zu", sizeof(dtm_train_state), sizeof(dtm_update_config));\n'
                   + "".join(f'  printf(" %zu", offsetof(dtm_train_state, {f}));\n' for f in fields_state)
                   + "".join(f'  printf(" %zu", offsetof(dtm_update_config, {f}));\n' for f in fields_cfg)
                   + "  return 0;\n}\n")
    exe = tmp_path / "layout"
    r = subprocess.run([cc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    want = [ctypes.sizeof(native.TrainState), ctypes.sizeof(native.UpdateConfig)]
    want += [getattr(native.TrainState, f).offset for f in fields_state]
    want += [getattr(native.UpdateConfig, f).offset for f in fields_cfg]
    assert got == want, (got, want)
    assert native.STATE_BYTES == got[0]


def test_plain_c_host_uses_the_library(tmp_path):
    """INTEGRATION.md's claim that the library is usable from a plain-C host: a C99 program built against include/dtm_b200.h
    checks the ABI version, computes a CRC-32C and reads batches from a TFRecord file through the loader and its batch ring -
    no Python, no GPU."""
    import shutil
    import subprocess

    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord
    from distributed_tensorflow_ibm_mnist_b200.ops import _lib

    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None or not _lib.available():
        pytest.skip("no C compiler or library not built")
    rec = str(tmp_path / "d.tfrecord")
    tfrecord.write_mnist_tfrecord(rec, [bytes([i]) * 784 for i in range(64)], [i % 10 for i in range(64)])
    src = tmp_path / "host.c"
    src.write_text(r'''
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "dtm_b200.h"
int main(int argc, char** argv) {
  if (dtm_abi_version() != DTM_ABI_VERSION) { printf("abi %d\n", dtm_abi_version()); return 1; }
  if (dtm_crc32c("123456789", 9, 0) != 0xE3069283u) return 2;
  const char* files[1] = {argv[1]};
  void* ld = dtm_loader_create(files, 1, 16, 784, 2, 7, /*shuffle=*/0, 0);
  if (!ld) { printf("create: %s\n", dtm_last_error()); return 3; }
  uint8_t* px = malloc(16 * 784); int32_t lab[16];
  if (dtm_loader_next(ld, px, lab) != 0) return 4;
  for (int i = 0; i < 16; ++i) if (px[i * 784] != i || lab[i] != i % 10) return 5;
  enum { NB = 3, BYTES = 16 * 784 + 16 * 4 };
  uint8_t* bufs[NB];
  for (int i = 0; i < NB; ++i) bufs[i] = malloc(BYTES);
  if (dtm_loader_ring_start(ld, bufs, NB, 16 * 784) != 0) { printf("ring: %s\n", dtm_last_error()); return 6; }
  for (int b = 1; b < 8; ++b) {                     /* batches 1..7: records 16 b .. 16 b + 15, wrapping after 64 */
    int idx = dtm_loader_ring_acquire(ld);
    if (idx < 0) return 7;
    const int32_t* l = (const int32_t*)(bufs[idx] + 16 * 784);
    for (int i = 0; i < 16; ++i) { int r = (16 * b + i) % 64; if (bufs[idx][i * 784] != r || l[i] != r % 10) return 8; }
    if (dtm_loader_ring_release(ld, idx) != 0) return 9;
  }
  if (dtm_loader_next(ld, px, lab) == 0 || !strstr(dtm_last_error(), "ring")) return 10;
  printf("records_read>=%d ok\n", (int)(dtm_loader_records_read(ld) >= 128));
  dtm_loader_destroy(ld);
  return 0;
}
''')
    exe = tmp_path / "host"
    libdir = str(_lib.lib_path().parent)
    r = subprocess.run([cc, "-std=c99", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src), "-L", libdir,
                        "-ldtm_b200", f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe), rec], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "records_read>=1 ok" in r.stdout, (r.returncode, r.stdout, r.stderr)


def test_native_loader_reports_damaged_and_empty_files(tmp_path):
    """Error paths of the C++ reader over mapped files: truncated header, truncated payload, corrupt length, empty input."""
    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord

    if readers._native_loader_lib() is None:
        pytest.skip("libdtm_b200.so not built")
    good = str(tmp_path / "good.tfrecord")
    tfrecord.write_mnist_tfrecord(good, [bytes([i]) * 784 for i in range(40)], [i % 10 for i in range(40)])
    raw = open(good, "rb").read()
    rec_len = len(raw) // 40

    def first_error(data, batch=8):
        p = str(tmp_path / "bad.tfrecord")
        open(p, "wb").write(data)
        ds = readers.TFRecordBatches([p], batch, 784, shuffle=False, native=True, prefetch_batches=0)
        try:
            for _ in range(12):
                ds.next()
        except ValueError as e:
            return str(e)
        finally:
            ds.close()
        return ""

    assert "truncated TFRecord header" in first_error(raw[:rec_len * 20 + 5])
    assert "truncated TFRecord in" in first_error(raw[:rec_len * 20 + 100])
    damaged = bytearray(raw)
    damaged[rec_len * 3 + 2] ^= 0x01                                        # inside the 4th record's length field
    assert "corrupt TFRecord length" in first_error(bytes(damaged))
    assert "no records" in first_error(b"")
    assert first_error(raw) == ""                                            # the undamaged file streams on across epochs
