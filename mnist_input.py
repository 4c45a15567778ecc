"""Model-definition module with the reference's function names (mnist_input.py: inputs / inference / loss / accuracy /
train), expressed eagerly over torch tensors.  The heavy lifting lives in the package; this module keeps scripts that
did ``import mnist_input`` working:

    images, labels = mnist_input.inputs(eval_data=False, batch_size=128)
    logits = mnist_input.inference(images)            # uses mnist_input.PARAMS (fresh reference-initialised variables)
    total  = mnist_input.loss(logits, labels)
    acc    = mnist_input.accuracy(logits, labels)
    step   = mnist_input.train(total, global_step)    # applies one optimizer step, returns the new global step
"""
from __future__ import annotations

import torch

import tf_parameter_mgr
from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets

batch_size = tf_parameter_mgr.getTrainBatchSize()

IMAGE_SIZE = presets.IMAGE_SIZE
NUM_CLASSES = presets.NUM_CLASSES
NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = presets.NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN
NUM_EXAMPLES_PER_EPOCH_FOR_EVAL = 10000
MOVING_AVERAGE_DECAY = presets.MOVING_AVERAGE_DECAY
NUM_EPOCHS_PER_DECAY = presets.NUM_EPOCHS_PER_DECAY
LEARNING_RATE_DECAY_FACTOR = tf_parameter_mgr.getLearningRateDecay()
INITIAL_LEARNING_RATE = tf_parameter_mgr.getBaseLearningRate()

SPEC = presets.make_spec(tf_parameter_mgr.params().model, tf_parameter_mgr.params().channels)
IMAGE_LAYERS = SPEC.channels
IMAGE_PIXELS = SPEC.image_pixels
PARAMS = {k: v.requires_grad_(True) for k, v in oracle.init_params(SPEC, seed=tf_parameter_mgr.params().seed).items()}
_OPT = None
_SOURCES: dict = {}


def inputs(eval_data, batch_size):
    """Normalised float images [B, IMAGE_PIXELS] and int32 labels [B] from the configured TFRecords (or synthetic data)."""
    files = tf_parameter_mgr.getTestData() if eval_data else tf_parameter_mgr.getTrainData()
    key = (bool(eval_data), batch_size)
    if key not in _SOURCES:
        if files:
            from distributed_tensorflow_ibm_mnist_b200.data.readers import TFRecordBatches

            _SOURCES[key] = TFRecordBatches(files, batch_size, IMAGE_PIXELS)
        else:
            from distributed_tensorflow_ibm_mnist_b200.data.synthetic import SyntheticMnist

            _SOURCES[key] = SyntheticMnist(batch_size, IMAGE_PIXELS, separable=True, pin=False)
    images, labels = _SOURCES[key].next()
    return oracle.normalise_u8(images), labels


def inference(images, params=None):
    return oracle.forward(params or PARAMS, images.reshape(images.shape[0], -1), SPEC)


def loss(logits, labels, params=None):
    return oracle.loss_fn(logits, labels, params or PARAMS, SPEC)[0]


def accuracy(logits, labels):
    return oracle.accuracy_fn(logits, labels)


def train(total_loss, global_step=None, return_grad=False):
    """One optimizer step on PARAMS from ``total_loss`` (staircase lr decay, optimizer from the parameter manager, EMA)."""
    global _OPT
    if _OPT is None:
        o = tf_parameter_mgr.getOptimizer(INITIAL_LEARNING_RATE)
        _OPT = oracle.ReferenceOptimizer(PARAMS, SPEC, optimizer=o.kind, base_lr=INITIAL_LEARNING_RATE,
                                         decay_rate=LEARNING_RATE_DECAY_FACTOR, decay_steps=presets.decay_steps(batch_size),
                                         momentum=o.momentum)
        if global_step is not None:
            _OPT.global_step = int(global_step)
    names = list(PARAMS)
    grads = dict(zip(names, torch.autograd.grad(total_loss, [PARAMS[n] for n in names])))
    for p in PARAMS.values():
        p.requires_grad_(False)
    _OPT.step(grads)
    for p in PARAMS.values():
        p.requires_grad_(True)
    if return_grad:
        return _OPT.global_step, grads
    return _OPT.global_step
