"""Worker of test_framework.py::test_dist_plane_passes_file_descriptors: two gloo ranks hand each other an open file."""
import os
import sys

import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_tensorflow_ibm_mnist_b200.parallel.control_plane import DistPlane  # noqa: E402


def main() -> int:
    rank, world, port, tmp = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4]
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    plane = DistPlane(dist.group.WORLD)
    path = os.path.join(tmp, f"rank{rank}.txt")
    with open(path, "w") as f:
        f.write(f"hello from rank {rank}")
    fd = os.open(path, os.O_RDONLY)
    got = plane.exchange_fds([fd])
    texts = []
    for j, fds in enumerate(got):
        assert len(fds) == 1
        texts.append(os.pread(fds[0], 64, 0).decode())
        if j != rank:
            os.close(fds[0])
    os.close(fd)
    assert texts == [f"hello from rank {j}" for j in range(world)], texts
    # a second round with descriptors from the root only (how the multicast handle travels)
    got = plane.exchange_fds([os.open(path, os.O_RDONLY)] if rank == 0 else [])
    assert [len(g) for g in got] == [1] + [0] * (world - 1)
    assert os.pread(got[0][0], 64, 0).decode() == "hello from rank 0"
    assert plane.all_true(True) and not plane.all_true(rank == 0) and plane.any_true(rank == 1) and not plane.any_true(False)
    assert plane.broadcast_object({"k": rank}, 1) == {"k": 1}
    dist.destroy_process_group()
    print("FD_EXCHANGE_OK", rank)
    return 0


if __name__ == "__main__":
    sys.exit(main())
