"""Worker of tests/test_framework.py::test_collective_saves_over_gloo: two ranks, an engine that claims to be sharded (like
the fused multi-GPU update), a saver whose clock only the chief consults.  Every rank must enter ``save`` the same number
of times, at the same steps, and only the chief may write files."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main() -> int:
    rank, world, port, out_dir = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4]
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing, trainer
    from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

    spec = presets.make_spec("baseline")
    eng = TorchEngine(spec, 8, optimizer="momentum", base_lr=0.01)
    eng.load_params(oracle.init_params(spec, seed=0))
    eng.sharded = True                                   # what NativeEngine reports with a PeerComm

    calls = []

    class Saver(checkpointing.CheckpointSaver):
        def due(self):                                   # only the chief's answer may matter
            return rank == 0 and trainer_loop.local_steps in (4, 8)

        def save(self, engine, write=True):
            calls.append((trainer_loop.local_steps, bool(write)))
            return super().save(engine, write=write)

    saver = Saver(os.path.join(out_dir, "train"), spec, "momentum", save_secs=1e9)
    src = SyntheticMnist(8, spec.image_pixels, pool_batches=2, seed=rank, pin=False)
    hooks = [trainer.StopAtStepHook(last_step=10)]
    trainer_loop = trainer.MonitoredTrainingLoop(eng, src, hooks=hooks, is_chief=rank == 0, saver=saver, task_id=rank,
                                                 group=dist.group.WORLD, save_poll_every=2, out=lambda *a: None)
    with trainer_loop as sess:
        while not sess.should_stop():
            sess.run_step()
    json.dump({"rank": rank, "calls": calls}, open(os.path.join(out_dir, f"calls{rank}.json"), "w"))
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
