"""Drop-in module for the reference's ``import tf_parameter_mgr`` (main.py:35, mnist_input.py:6, inference.py:12)."""
from distributed_tensorflow_ibm_mnist_b200.config.parameter_mgr import *  # noqa: F401,F403
from distributed_tensorflow_ibm_mnist_b200.config.parameter_mgr import (  # noqa: F401
    configure, getBaseLearningRate, getLearningRateDecay, getMaxSteps, getOptimizer, getTestData, getTestInterval,
    getTrainBatchSize, getTrainData, getValData, load_file, params, reset)
