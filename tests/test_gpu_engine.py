"""Whole-model checks of the native engine against the PyTorch oracle (same weights, same batch, same dropout mask)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(preset, B, seed=0, **kw):
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    spec = presets.make_spec(preset)
    params = oracle.init_params(spec, seed=seed)
    # larger-than-default head weights so that the loss gradient is not vanishingly small
    eng = NativeEngine(spec, B, seed=1234, use_graph=kw.pop("use_graph", False), **kw)
    eng.load_params(params)
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randint(0, 256, (B, spec.image_pixels), dtype=torch.uint8, generator=g)
    y = torch.randint(0, 10, (B,), dtype=torch.int32, generator=g)
    eng.stage_batch(0, x.cuda(), y.cuda())
    return spec, params, eng, x, y


def _oracle_grads(spec, params, x, y, step, seed=1234):
    from distributed_tensorflow_ibm_mnist_b200.models import oracle
    from distributed_tensorflow_ibm_mnist_b200.ops.philox import dropout_keep_mask

    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    mask = None
    if spec.keep_prob < 1.0:
        mask = torch.from_numpy(dropout_keep_mask(x.shape[0], spec.head_in, seed=seed, step=step, keep_prob=spec.keep_prob))
    logits = oracle.forward(p, oracle.normalise_u8(x), spec, dropout_mask=mask, emulate_bf16=True)
    total, xent = oracle.loss_fn(logits, y, p, spec)
    xent.backward()  # weight-decay gradients are added by the update kernel, not by backward
    return logits.detach(), xent.item(), {k: v.grad for k, v in p.items()}


def _rel(a, b):
    return (a.float() - b.float()).abs().max().item() / (b.float().abs().max().item() + 1e-20)


@pytest.mark.parametrize("preset", ["baseline", "reference"])
def test_forward_backward_matches_oracle(preset):
    B = 64
    spec, params, eng, x, y = _setup(preset, B)
    eng.forward_backward(0)
    torch.cuda.synchronize()
    logits, xent, grads = _oracle_grads(spec, params, x, y, step=0)
    assert _rel(eng.logits.cpu(), logits) < 3e-2
    from distributed_tensorflow_ibm_mnist_b200.ops import native

    st = native.read_state(eng.eval_state)
    assert abs(st.loss_sum - xent) < 2e-2 * max(1.0, abs(xent))
    got = eng.gradients()
    # The oracle rounds activations AND the stored gradients (dP2, dZ2, dH, pooled conv1 gradient) to bf16 where the kernels do.
    # Measured on a B200 over three seeds (relative L2 / worst element relative to the tensor's largest): baseline preset
    # <= 0.006 / 0.028; reference preset (LRN twice, a second fc layer: single ReLU gates flip between the bf16 and the fp32
    # pre-activations) <= 0.048 / 0.115.
    l2_tol, max_tol = (2e-2, 6e-2) if preset == "baseline" else (6e-2, 0.15)
    for k, g in grads.items():
        l2 = ((got[k].float() - g).norm() / (g.norm() + 1e-20)).item()
        assert l2 < l2_tol, (k, l2)
        assert _rel(got[k], g) < max_tol, (k, _rel(got[k], g))


@pytest.mark.parametrize("use_graph", [False, True])
def test_training_reduces_loss_and_tracks_oracle(use_graph):
    from distributed_tensorflow_ibm_mnist_b200.models import oracle

    B = 64
    # momentum, not Adam: Adam's sign-like update turns rounding noise on near-zero gradients into O(lr) weight
    # differences, which makes an element-wise weight comparison meaningless
    spec, params, eng, x, y = _setup("baseline", B, use_graph=use_graph, optimizer="momentum", base_lr=0.02)
    ref_params = {k: v.clone() for k, v in params.items()}
    opt = oracle.ReferenceOptimizer(ref_params, spec, optimizer="momentum", base_lr=0.02)
    losses, ref_losses = [], []
    for step in range(6):
        eng.train_step(0)
        m = eng.metrics()
        losses.append(m["cross_entropy"])
        assert m["global_step"] == step + 1 and not m["nan"]
        _, xent, grads = _oracle_grads(spec, ref_params, x, y, step=step)
        ref_losses.append(xent)
        opt.step(grads)
    assert losses[-1] < losses[0]          # same batch every step: the loss must go down
    for a, b in zip(losses, ref_losses):
        assert abs(a - b) < 5e-2 * max(1.0, abs(b)), (losses, ref_losses)
    got = eng.params()
    for k in ("conv2/weights", "local3/weights", "softmax_linear/weights"):
        assert _rel(got[k], ref_params[k]) < 5e-2, k


def test_adam_training_tracks_oracle_loss():
    from distributed_tensorflow_ibm_mnist_b200.models import oracle

    spec, params, eng, x, y = _setup("baseline", 64, use_graph=True, optimizer="adam", base_lr=1e-3)
    ref_params = {k: v.clone() for k, v in params.items()}
    opt = oracle.ReferenceOptimizer(ref_params, spec, optimizer="adam", base_lr=1e-3)
    for step in range(5):
        eng.train_step(0)
        got = eng.metrics()["cross_entropy"]
        _, xent, grads = _oracle_grads(spec, ref_params, x, y, step=step)
        opt.step(grads)
        assert abs(got - xent) < 5e-2 * max(1.0, abs(xent)), (step, got, xent)


def test_predict_matches_oracle_probabilities():
    from distributed_tensorflow_ibm_mnist_b200.models import oracle

    spec, params, eng, x, y = _setup("baseline", 32)
    probs = eng.predict(x[:20].cuda()).cpu()
    want = torch.softmax(oracle.forward(params, oracle.normalise_u8(x[:20]), spec, emulate_bf16=True), 1)
    assert (probs - want).abs().max().item() < 2e-2
    assert abs(probs.sum(1) - 1).max().item() < 1e-4
