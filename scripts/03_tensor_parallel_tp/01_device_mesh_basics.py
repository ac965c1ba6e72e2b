#!/usr/bin/env python
"""DeviceMesh basics: 1-D ("tp",) mesh, 2-D (dp, tp) mesh with tp = min(4, world), named slicing and an all-reduce
sanity check that is ASSERTED, not just printed (/root/reference/scripts/03_tensor_parallel_tp/01_device_mesh_basics.py)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from _common import setup, torch  # noqa: E402


def main():
    rank, world, local, device = setup()
    from b200dist.parallel import init_device_mesh
    from b200dist.utils import cleanup_distributed
    from b200dist.utils.distributed import barrier

    mesh1 = init_device_mesh(device.type, (world,), ("tp",))
    print(f"[rank {rank}] 1-D mesh: {mesh1}", flush=True)
    barrier()
    tp = min(4, world)
    while world % tp:
        tp -= 1
    mesh2 = init_device_mesh(device.type, (world // tp, tp), ("dp", "tp"))
    tp_mesh, dp_mesh = mesh2["tp"], mesh2["dp"]
    print(f"[rank {rank}] 2-D mesh {mesh2.shape}: dp rank {mesh2.get_local_rank('dp')} tp rank {mesh2.get_local_rank('tp')} "
          f"tp group {tp_mesh.ranks()} dp group {dp_mesh.ranks()}", flush=True)
    barrier()
    t = torch.tensor([float(rank)] * 8, device=device)
    tp_mesh.communicator().all_reduce(t)
    expect = float(sum(tp_mesh.ranks()))
    assert torch.all(t == expect), (t, expect)
    if rank == 0:
        print(f"all-reduce over the tp group: {t[0].item()} == sum{tp_mesh.ranks()} ✓", flush=True)
    barrier()
    cleanup_distributed()


if __name__ == "__main__":
    main()
