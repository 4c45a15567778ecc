"""Drop-in module for the reference's ``import inference_helper`` (inference.py:12)."""
from distributed_tensorflow_ibm_mnist_b200.utils.inference_helper import writeClassificationResult  # noqa: F401
