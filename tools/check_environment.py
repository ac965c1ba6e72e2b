#!/usr/bin/env python
"""Environment report for one NVSwitch box (capability twin of /root/reference/tests/check_environment.py): ranks -> GPUs,
PyTorch / CUDA / cuDNN / NCCL versions, NCCL env vars, NVLink peer access matrix, the native library and its sm_100a
kernels, a peer-memory all-reduce sanity check, and a ✓/✗ summary that — unlike the reference (:351-373) — is reflected in
the exit code."""
from __future__ import annotations

import os
import socket
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

NCCL_VARS = ["NCCL_DEBUG", "NCCL_DEBUG_SUBSYS", "NCCL_SOCKET_IFNAME", "NCCL_IB_DISABLE", "NCCL_SHM_DISABLE", "NCCL_P2P_DISABLE",
             "NCCL_NVLS_ENABLE", "NCCL_ALGO", "NCCL_PROTO", "NCCL_TIMEOUT", "NCCL_ASYNC_ERROR_HANDLING", "NCCL_BLOCKING_WAIT",
             "TORCH_DISTRIBUTED_DEBUG", "CUDA_VISIBLE_DEVICES", "B200DIST_STAGING_MB"]


def main():
    from b200dist.utils import cleanup_distributed, init_distributed
    from b200dist.utils.distributed import launcher_name

    rank, world, local = init_distributed(verbose=False)
    checks = {}
    cuda = torch.cuda.is_available()
    info = {"rank": rank, "host": socket.gethostname(), "local_rank": local,
            "gpu": torch.cuda.get_device_name() if cuda else "cpu",
            "mem_gb": round(torch.cuda.get_device_properties(0).total_memory / 2**30, 1) if cuda else 0,
            "cc": ".".join(map(str, torch.cuda.get_device_capability())) if cuda else "-"}
    infos = [None] * world
    if world > 1:
        dist.all_gather_object(infos, info)
    else:
        infos = [info]
    if rank == 0:
        print(f"launcher: {launcher_name()}  world size: {world}")
        print(f"PyTorch {torch.__version__}  CUDA {torch.version.cuda}  cuDNN {torch.backends.cudnn.version() if cuda else '-'}  "
              f"NCCL {'.'.join(map(str, torch.cuda.nccl.version())) if cuda else '-'}")
        print("rank -> device:")
        for i in infos:
            print(f"  rank {i['rank']:>2}  {i['host']}  local {i['local_rank']}  {i['gpu']}  {i['mem_gb']} GB  sm_{i['cc'].replace('.', '')}")
        print("environment:")
        for k in NCCL_VARS:
            if k in os.environ:
                print(f"  {k}={os.environ[k]}")
        if cuda:
            n = torch.cuda.device_count()
            print("peer access (NVLink / NVSwitch):")
            for i in range(n):
                print("  " + " ".join("✓" if i == j or torch.cuda.can_device_access_peer(i, j) else "✗" for j in range(n)))
    checks["cuda available"] = cuda
    checks["sm_100 (Blackwell)"] = cuda and torch.cuda.get_device_capability()[0] == 10
    try:
        from b200dist.ops import _lib

        lib = _lib.load()
        checks["native library loads"] = lib.b200_version() >= 100
    except Exception as e:  # noqa: BLE001
        checks["native library loads"] = False
        print(f"[rank {rank}] native library: {e}")
    if cuda and checks["native library loads"]:
        from b200dist.comm import get_communicator
        from b200dist.ops import native as N

        a = torch.randn(256, 256, device="cuda").bfloat16()
        checks["tcgen05 GEMM"] = bool(torch.allclose(N.gemm(a, a).float(), a.float() @ a.float().t(), atol=0.5, rtol=0.05))
        t = torch.full((1024,), float(rank), device="cuda")
        get_communicator(None).all_reduce(t)
        checks["peer-memory all-reduce"] = bool((t == sum(range(world))).all())
        try:   # informational: NVLS needs multicast support on every GPU (the pull kernels are used otherwise)
            from b200dist.comm.peer_mem import multicast_supported

            if rank == 0:
                print(f"NVSwitch multicast (NVLS all-reduce path): {'available' if multicast_supported() else 'not available'}")
        except Exception as e:  # noqa: BLE001
            print(f"[rank {rank}] multicast query failed: {e}")
    ok = all(checks.values())
    if rank == 0:
        print("summary:")
        for k, v in checks.items():
            print(f"  {'✓' if v else '✗'} {k}")
    flag = torch.tensor([0 if ok else 1], device="cuda" if cuda else "cpu")
    if world > 1:
        dist.all_reduce(flag)
    cleanup_distributed()
    sys.exit(0 if int(flag) == 0 else 1)


if __name__ == "__main__":
    main()
