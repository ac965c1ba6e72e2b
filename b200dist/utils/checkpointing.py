"""Checkpoint save / load.

Two on-disk layouts:

1. **Consolidated** — byte-compatible with the reference helper
   (/root/reference/utils/checkpointing.py:23-88): ONE file written by rank 0 with keys
   ``model_state_dict``, ``optimizer_state_dict``, ``epoch``; ``.module`` prefixes of wrappers are
   stripped; every rank then meets at a barrier.  A file written here loads with the reference's
   ``load_checkpoint`` and vice versa.
2. **Sharded** — a directory with ``index.json`` + one ``shard_{rank:05d}.pt`` per rank, for
   FSDP/TP-sharded models whose full state does not fit (or should not be gathered) on one
   rank.  ``consolidate_sharded_checkpoint`` turns it into layout 1, so the sharded layout
   always round-trips to the reference format.

Also covered: the DDP-basic snapshot format ``{MODEL_STATE, EPOCHS_RUN}``
(/root/reference/scripts/01_data_parallel_ddp/multinode_ddp_basic.py:150-155,184-196) and the
FSDP "gather full state dict to rank-0 CPU" path
(/root/reference/scripts/02_fully_sharded_fsdp/multinode_fsdp_unet.py:285-298).
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist


def _is_dist() -> bool:
    return dist.is_available() and dist.is_initialized()


def _rank() -> int:
    return dist.get_rank() if _is_dist() else 0


def _barrier() -> None:
    if _is_dist():
        if dist.get_backend() == "nccl" and torch.cuda.is_available():
            dist.barrier(device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier()


def unwrap_model(model):
    """Strip parallel wrappers (ours expose ``.module`` exactly like DDP does)."""
    while hasattr(model, "module") and isinstance(getattr(model, "module"), torch.nn.Module):
        model = model.module
    return model


def full_state_dict(model, rank0_only: bool = True, offload_to_cpu: bool = True) -> Dict[str, torch.Tensor]:
    """Unsharded state dict with reference parameter names.

    Wrappers that shard parameters (FSDP, TP) implement ``full_state_dict()`` — a collective
    that all ranks must enter; rank 0 receives the tensors (others get ``{}`` when
    ``rank0_only``).  Plain / DDP modules just return ``state_dict()``.
    """
    m = model
    while m is not None:
        fn = getattr(m, "full_state_dict", None)
        if callable(fn) and not isinstance(m, type):
            return fn(rank0_only=rank0_only, offload_to_cpu=offload_to_cpu)
        m = getattr(m, "module", None) if isinstance(getattr(m, "module", None), torch.nn.Module) else None
    sd = unwrap_model(model).state_dict()
    if offload_to_cpu:
        sd = {k: v.detach().cpu() for k, v in sd.items()}
    return sd


def save_checkpoint(model, optimizer, epoch: int, path: str, rank: Optional[int] = None) -> None:
    """Rank-0 single-file checkpoint in the reference format, followed by a barrier."""
    if rank is None:
        rank = _rank()
    state = full_state_dict(model)  # collective for sharded wrappers: all ranks enter
    # a sharded optimizer (FSDP) gathers its fp32 master / moments into unsharded tensors - also a collective - so the
    # single file never holds ONE rank's shards (which every rank would then load as its own)
    gather_opt = getattr(optimizer, "full_state_dict", None)
    opt_full = gather_opt() if callable(gather_opt) else None
    if rank == 0:
        d = os.path.dirname(os.path.abspath(path))
        os.makedirs(d, exist_ok=True)
        opt_state = opt_full if opt_full is not None else (optimizer.state_dict() if optimizer is not None else {})
        tmp = path + ".tmp"
        torch.save({"model_state_dict": state, "optimizer_state_dict": opt_state, "epoch": epoch}, tmp)
        os.replace(tmp, path)
    _barrier()


def load_checkpoint(model, optimizer, path: str, device=None) -> int:
    """Restore model (+ optimizer) and return the stored epoch; 0 when the file is missing."""
    if not os.path.exists(path):
        return 0
    ckpt = torch.load(path, map_location=device or "cpu", weights_only=False)
    target = model
    loader = getattr(model, "load_full_state_dict", None)
    if callable(loader):
        loader(ckpt["model_state_dict"])
    else:
        target = unwrap_model(model)
        target.load_state_dict(ckpt["model_state_dict"])
    if optimizer is not None and ckpt.get("optimizer_state_dict"):
        try:
            optimizer.load_state_dict(ckpt["optimizer_state_dict"])
        except (ValueError, KeyError):
            pass  # optimizer layout differs (e.g. sharded vs replicated): weights still restored
    return int(ckpt.get("epoch", 0))


# ----------------------------- DDP-basic snapshot ---------------------------------
def save_snapshot(model, epochs_run: int, path: str = "snapshot.pt") -> None:
    """``{MODEL_STATE, EPOCHS_RUN}``, written ONCE per job (global rank 0), not once per node as
    the reference does (multinode_ddp_basic.py:195, SURVEY.md §9 defects)."""
    state = full_state_dict(model)
    if _rank() == 0:
        tmp = path + ".tmp"
        torch.save({"MODEL_STATE": state, "EPOCHS_RUN": epochs_run}, tmp)
        os.replace(tmp, path)
    _barrier()


def load_snapshot(model, path: str = "snapshot.pt", device=None) -> int:
    if not os.path.exists(path):
        return 0
    snap = torch.load(path, map_location=device or "cpu", weights_only=False)
    unwrap_model(model).load_state_dict(snap["MODEL_STATE"])
    return int(snap.get("EPOCHS_RUN", 0))


# ----------------------------- sharded layout --------------------------------------
def save_sharded_checkpoint(model, optimizer, epoch: int, directory: str) -> None:
    """Every rank writes its local shard; rank 0 writes ``index.json``.

    The model must provide ``sharded_state_dict() -> {name: ShardEntry}`` where each entry is a
    dict ``{"tensor": local_tensor, "global_shape": [...], "dim": int|None, "index": i, "parts": n,
    "flat_range": [start, stop]|None}``; plain modules are treated as fully replicated
    (rank 0 stores them, ``parts == 1``).
    """
    rank = _rank()
    world = dist.get_world_size() if _is_dist() else 1
    os.makedirs(directory, exist_ok=True)
    fn = getattr(model, "sharded_state_dict", None)
    if callable(fn):
        entries = fn()
    else:
        entries = {k: {"tensor": v, "global_shape": list(v.shape), "dim": None, "index": 0, "parts": 1,
                       "flat_range": None} for k, v in unwrap_model(model).state_dict().items()}
        if rank != 0:
            entries = {}
    payload = {"model": {k: {**e, "tensor": e["tensor"].detach().cpu()} for k, e in entries.items()},
               "optimizer": _to_cpu(optimizer.state_dict()) if optimizer is not None else {}, "epoch": epoch,
               # per-rank generator state: an exact resume replays the same dropout masks / synthetic batches
               "rng": {"cpu": torch.get_rng_state(),
                       "cuda": torch.cuda.get_rng_state() if torch.cuda.is_available() else None}}
    torch.save(payload, os.path.join(directory, f"shard_{rank:05d}.pt"))
    if rank == 0:
        with open(os.path.join(directory, "index.json"), "w") as f:
            json.dump({"format": "b200dist-sharded-v1", "world_size": world, "epoch": epoch}, f)
    _barrier()


def _to_cpu(obj):
    if torch.is_tensor(obj):
        return obj.detach().cpu()
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def _assemble(name: str, pieces: list) -> torch.Tensor:
    e0 = pieces[0]
    if e0["parts"] == 1 or (e0["dim"] is None and e0["flat_range"] is None):
        return e0["tensor"]
    if e0["flat_range"] is not None:  # FSDP flat shards: contiguous ranges of the flattened tensor
        numel = 1
        for s in e0["global_shape"]:
            numel *= s
        flat = torch.empty(numel, dtype=e0["tensor"].dtype)
        for e in pieces:
            a, b = e["flat_range"]
            flat[a:b] = e["tensor"].reshape(-1)[: b - a]
        return flat.view(e0["global_shape"])
    pieces = sorted(pieces, key=lambda e: e["index"])
    return torch.cat([e["tensor"] for e in pieces], dim=e0["dim"])


def load_sharded_state(directory: str) -> Dict[str, Any]:
    """Read every shard and assemble the full (unsharded) state on the calling process."""
    with open(os.path.join(directory, "index.json")) as f:
        index = json.load(f)
    by_name: Dict[str, list] = {}
    opt_state, epoch = {}, index.get("epoch", 0)
    for r in range(index["world_size"]):
        shard = torch.load(os.path.join(directory, f"shard_{r:05d}.pt"), map_location="cpu", weights_only=False)
        for k, e in shard["model"].items():
            by_name.setdefault(k, []).append(e)
        if r == 0:
            opt_state = shard.get("optimizer", {})
    model_state = {k: _assemble(k, v) for k, v in by_name.items()}
    return {"model_state_dict": model_state, "optimizer_state_dict": opt_state, "epoch": epoch}


def consolidate_sharded_checkpoint(directory: str, out_path: str) -> None:
    """Sharded directory -> single reference-format file (layout 1)."""
    torch.save(load_sharded_state(directory), out_path)


def load_sharded_checkpoint(model, optimizer, directory: str) -> int:
    if not os.path.exists(os.path.join(directory, "index.json")):
        return 0
    full = load_sharded_state(directory)
    loader = getattr(model, "load_full_state_dict", None)
    if callable(loader):
        loader(full["model_state_dict"])
    else:
        unwrap_model(model).load_state_dict(full["model_state_dict"])
    # optimizer + generator state are per rank: restored when the job has the world size the checkpoint was written with
    with open(os.path.join(directory, "index.json")) as f:
        saved_world = json.load(f)["world_size"]
    world = dist.get_world_size() if _is_dist() else 1
    mine = os.path.join(directory, f"shard_{_rank():05d}.pt")
    if saved_world == world and os.path.exists(mine):
        shard = torch.load(mine, map_location="cpu", weights_only=False)
        if optimizer is not None and shard.get("optimizer"):
            optimizer.load_state_dict(shard["optimizer"])
        rng = shard.get("rng") or {}
        if rng.get("cpu") is not None:
            torch.set_rng_state(rng["cpu"])
        if rng.get("cuda") is not None and torch.cuda.is_available():
            torch.cuda.set_rng_state(rng["cuda"])
    return int(full["epoch"])
