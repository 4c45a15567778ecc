"""``graph.pbtxt`` and ``model.ckpt-N.meta`` — the two graph artefacts a TensorFlow training directory carries.

The reference's ``MonitoredTrainingSession`` (main.py:140-146) writes ``<train_dir>/graph.pbtxt`` when the chief starts and
a ``MetaGraphDef`` next to every checkpoint (``Saver(write_meta_graph=True)`` default).  This framework has no TensorFlow
graph, but tools that browse a training directory (TensorBoard's graph tab, ``inspect_checkpoint``-style scripts,
``tf.train.import_meta_graph`` for the variable list) expect the files, so the chief writes an honest, minimal version:
one ``VariableV2`` node per checkpointed variable with its dtype and shape, plus the ``SaverDef`` describing the bundle
(V2 format, sharded, ``max_to_keep``).  Hand-encoded protobuf like the rest of ``formats/`` (no TensorFlow, no generated
code); field numbers are TensorFlow's (graph.proto, node_def.proto, attr_value.proto, tensor_shape.proto,
meta_graph.proto, saver.proto) and are checked in the tests against ``tensorboard.compat.proto``.
"""
from __future__ import annotations

import os

import numpy as np

from .proto import f_bytes, f_float, f_str, f_varint

DT_FLOAT, DT_INT32, DT_INT64 = 1, 3, 9
_DTYPES = {np.dtype(np.float32): DT_FLOAT, np.dtype(np.int32): DT_INT32, np.dtype(np.int64): DT_INT64}
_DT_NAMES = {DT_FLOAT: "DT_FLOAT", DT_INT32: "DT_INT32", DT_INT64: "DT_INT64"}
GRAPH_PRODUCER = 26          # GraphDef version of the TF-1.x line the reference targets
FRAMEWORK_VERSION = "b200-dist-mnist-1"


def _dtype_enum(dtype) -> int:
    dt = np.dtype(dtype)
    if dt not in _DTYPES:
        raise ValueError(f"unsupported variable dtype {dt}")
    return _DTYPES[dt]


def _shape_proto(shape) -> bytes:
    return b"".join(f_bytes(2, f_varint(1, int(d))) for d in shape)          # TensorShapeProto.dim{size}


def _attr(key: str, value: bytes) -> bytes:
    return f_bytes(5, f_str(1, key) + f_bytes(2, value))                      # NodeDef.attr map entry


def variable_node(name: str, dtype, shape) -> bytes:
    """NodeDef of a ``VariableV2`` op (attrs: container, dtype, shape, shared_name — alphabetical like TF emits them)."""
    body = f_str(1, name) + f_str(2, "VariableV2")
    body += _attr("container", f_bytes(2, b""))
    body += _attr("dtype", f_varint(6, _dtype_enum(dtype)))
    body += _attr("shape", f_bytes(7, _shape_proto(shape)))
    body += _attr("shared_name", f_bytes(2, b""))
    return body


def graph_def(variables: dict[str, tuple]) -> bytes:
    """Binary GraphDef with one VariableV2 node per ``name -> (dtype, shape)`` (insertion order kept)."""
    out = b"".join(f_bytes(1, variable_node(n, dt, sh)) for n, (dt, sh) in variables.items())
    return out + f_bytes(4, f_varint(1, GRAPH_PRODUCER))                      # versions{producer}


def graph_pbtxt(variables: dict[str, tuple]) -> str:
    """The same graph in protobuf text format (what ``tf.train.write_graph(..., 'graph.pbtxt')`` produces)."""
    lines: list[str] = []
    for name, (dtype, shape) in variables.items():
        lines += ["node {", f'  name: "{name}"', '  op: "VariableV2"',
                  "  attr {", '    key: "container"', "    value {", '      s: ""', "    }", "  }",
                  "  attr {", '    key: "dtype"', "    value {", f"      type: {_DT_NAMES[_dtype_enum(dtype)]}", "    }", "  }",
                  "  attr {", '    key: "shape"', "    value {", "      shape {"]
        for d in shape:
            lines += ["        dim {", f"          size: {int(d)}", "        }"]
        lines += ["      }", "    }", "  }",
                  "  attr {", '    key: "shared_name"', "    value {", '      s: ""', "    }", "  }", "}"]
    lines += ["versions {", f"  producer: {GRAPH_PRODUCER}", "}"]
    return "\n".join(lines) + "\n"


def saver_def(max_to_keep: int = 5, sharded: bool = True) -> bytes:
    """SaverDef of the default Scaffold saver: V2 bundles, sharded, ``keep_checkpoint_every_n_hours`` = 10000."""
    return (f_str(1, "save/Const:0") + f_str(2, "save/Identity:0") + f_str(3, "save/restore_all") + f_varint(4, max_to_keep)
            + f_varint(5, 1 if sharded else 0) + f_float(6, 10000.0) + f_varint(7, 2))


def meta_graph_def(variables: dict[str, tuple], *, max_to_keep: int = 5, sharded: bool = True) -> bytes:
    meta_info = f_str(5, FRAMEWORK_VERSION) + f_str(6, "hand-encoded; no TensorFlow graph behind it")
    return f_bytes(1, meta_info) + f_bytes(2, graph_def(variables)) + f_bytes(3, saver_def(max_to_keep, sharded))


def describe(arrays: dict[str, np.ndarray]) -> dict[str, tuple]:
    """``name -> (dtype, shape)`` of checkpoint contents (the argument of the functions above)."""
    return {k: (np.asarray(v).dtype, tuple(np.asarray(v).shape)) for k, v in arrays.items()}


def write_graph_pbtxt(train_dir: str, variables: dict[str, tuple]) -> str:
    os.makedirs(train_dir, exist_ok=True)
    path = os.path.join(train_dir, "graph.pbtxt")
    tmp = path + ".tmp"
    with open(tmp, "w") as f:
        f.write(graph_pbtxt(variables))
    os.replace(tmp, path)
    return path


def write_meta(prefix: str, variables: dict[str, tuple], *, max_to_keep: int = 5, sharded: bool = True) -> str:
    path = prefix + ".meta"
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(meta_graph_def(variables, max_to_keep=max_to_keep, sharded=sharded))
    os.replace(tmp, path)
    return path
