"""The ``checkpoint`` state file that sits next to TensorFlow checkpoints.

``tf.train.get_checkpoint_state(dir)`` (reference inference.py:89) parses this text-format ``CheckpointState`` proto
to find the newest checkpoint; ``MonitoredTrainingSession`` (main.py:140-146) rewrites it after every save and keeps
the last ``max_to_keep`` (5) checkpoints.
"""
from __future__ import annotations

import os
import re
from dataclasses import dataclass, field

STATE_FILE = "checkpoint"


@dataclass
class CheckpointState:
    model_checkpoint_path: str = ""
    all_model_checkpoint_paths: list[str] = field(default_factory=list)


def _unescape(s: str) -> str:
    return bytes(s, "utf-8").decode("unicode_escape")


def _escape(s: str) -> str:
    return s.replace("\\", "\\\\").replace('"', '\\"')


def read_state(directory: str) -> CheckpointState | None:
    path = os.path.join(directory, STATE_FILE)
    if not os.path.exists(path):
        return None
    st = CheckpointState()
    with open(path, "r", encoding="utf-8") as f:
        for line in f:
            m = re.match(r'\s*(\w+)\s*:\s*"((?:[^"\\]|\\.)*)"', line)
            if not m:
                continue
            key, val = m.group(1), _unescape(m.group(2))
            if key == "model_checkpoint_path":
                st.model_checkpoint_path = val
            elif key == "all_model_checkpoint_paths":
                st.all_model_checkpoint_paths.append(val)
    return st


def write_state(directory: str, st: CheckpointState) -> None:
    os.makedirs(directory, exist_ok=True)
    lines = [f'model_checkpoint_path: "{_escape(st.model_checkpoint_path)}"']
    lines += [f'all_model_checkpoint_paths: "{_escape(p)}"' for p in st.all_model_checkpoint_paths]
    tmp = os.path.join(directory, f".{STATE_FILE}.tmp{os.getpid()}")
    with open(tmp, "w", encoding="utf-8") as f:
        f.write("\n".join(lines) + "\n")
    os.replace(tmp, os.path.join(directory, STATE_FILE))


def latest_checkpoint(directory: str) -> str | None:
    """Absolute prefix of the newest checkpoint in ``directory`` or None (like ``tf.train.latest_checkpoint``)."""
    st = read_state(directory)
    if st is None or not st.model_checkpoint_path:
        return None
    p = st.model_checkpoint_path
    if not os.path.isabs(p):
        p = os.path.join(directory, p)
    return p if os.path.exists(p + ".index") else None
