"""Drop-in module for the reference's ``import monitor_cb`` (main.py:36): CMonitor and the two collection keys."""
from distributed_tensorflow_ibm_mnist_b200.train.monitor import CMonitor, DLMAO_TEST_SUMMARIES, DLMAO_TRAIN_SUMMARIES  # noqa: F401
