"""Runtime utilities on CPU: launcher detection, config, logging, redirect, sampler, checkpoints."""
import os
import sys
from pathlib import Path

import pytest
import torch
import torch.nn as nn

from tests.dist_utils import run_distributed

REF = Path("/root/reference")


def _clean_env(monkeypatch):
    for k in ["RANK", "LOCAL_RANK", "WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK", "OMPI_COMM_WORLD_SIZE",
              "OMPI_COMM_WORLD_RANK", "PMI_RANK", "PMI_SIZE", "PMI_LOCAL_RANK", "PALS_LOCAL_RANKID", "MASTER_ADDR",
              "MASTER_PORT"]:
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("B200DIST_DISABLE_MPI4PY", "1")


def test_launcher_detection_priority(monkeypatch):
    from b200dist.utils.distributed import get_rank_info

    _clean_env(monkeypatch)
    assert get_rank_info() == (0, 1, 0, "single")
    monkeypatch.setenv("PMI_RANK", "5"); monkeypatch.setenv("PMI_SIZE", "8"); monkeypatch.setenv("PALS_LOCAL_RANKID", "1")
    assert get_rank_info() == (1, 8, 5, "cray-mpich")
    assert os.environ["MASTER_PORT"] == "29500" and "MASTER_ADDR" in os.environ
    monkeypatch.setenv("PMI_LOCAL_RANK", "3")
    assert get_rank_info()[0] == 3  # PMI_LOCAL_RANK wins over PALS_LOCAL_RANKID
    monkeypatch.setenv("OMPI_COMM_WORLD_LOCAL_RANK", "2"); monkeypatch.setenv("OMPI_COMM_WORLD_SIZE", "4")
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "2")
    assert get_rank_info() == (2, 4, 2, "openmpi")  # Open MPI before PMI
    monkeypatch.setenv("RANK", "7"); monkeypatch.setenv("LOCAL_RANK", "3"); monkeypatch.setenv("WORLD_SIZE", "16")
    assert get_rank_info() == (3, 16, 7, "torchrun")  # torchrun first


def test_init_single_process_is_noop(monkeypatch):
    from b200dist.utils import cleanup_distributed, init_distributed, is_main_rank

    _clean_env(monkeypatch)
    assert init_distributed(backend="nccl", verbose=False) == (0, 1, 0)  # CPU box: nccl -> gloo, no PG for world 1
    assert is_main_rank()
    cleanup_distributed()


def _init_worker(rank, world):
    import torch.distributed as dist

    from b200dist.utils import init_distributed, print_rank0

    # our own init path (env:// from the torchrun-style variables); the harness did not create a process group
    r, w, l = init_distributed(backend="nccl", verbose=False)
    assert (r, w, l) == (rank, world, rank)
    t = torch.tensor([float(rank)])
    dist.all_reduce(t)
    assert t.item() == sum(range(world))
    print_rank0("ok")


def test_init_distributed_gloo_world2():
    run_distributed(_init_worker, 2, init_pg=False)


def test_training_config_args_yaml(tmp_path):
    from b200dist.utils import TrainingConfig

    c = TrainingConfig.from_args(["--lr", "0.5", "--batch_size", "64", "--unknown-flag", "1", "--tp", "4", "--use_amp"])
    assert c.learning_rate == 0.5 and c.batch_size == 64 and c.tp == 4 and c.use_amp
    y = tmp_path / "c.yaml"
    y.write_text("num_epochs: 3\nlearning_rate: 0.01\nnot_a_field: 9\nbackend: gloo\n")
    c = TrainingConfig.from_yaml(str(y))
    assert c.num_epochs == 3 and c.learning_rate == 0.01 and c.backend == "gloo"
    c = TrainingConfig.from_args(["--config", str(y), "--num_epochs", "5"])
    assert c.num_epochs == 5 and c.backend == "gloo"
    ref_fields = ["num_epochs", "batch_size", "learning_rate", "weight_decay", "momentum", "seed", "backend", "use_amp",
                  "amp_dtype", "save_every", "checkpoint_dir", "profile", "profile_dir"]
    assert all(f in c.to_dict() for f in ref_fields)


def test_redirect_captures_fd_level_output(tmp_path, monkeypatch):
    from b200dist.utils import redirect

    monkeypatch.setenv("RANK", "3")
    with redirect(log_prefix="log", path=str(tmp_path)):
        print("python-level")
        os.write(1, b"fd-level\n")
        os.write(2, b"err-level\n")
    out = (tmp_path / "log.3.out").read_text()
    assert "python-level" in out and "fd-level" in out
    assert "err-level" in (tmp_path / "log.3.err").read_text()


def test_rank_log_only_rank0(caplog):
    import logging

    from b200dist.utils import get_logger, rank_log

    lg = get_logger("t")
    with caplog.at_level(logging.INFO):
        rank_log(1, lg, "hidden")
        rank_log(0, lg, "shown")
    assert "shown" in caplog.text and "hidden" not in caplog.text


def test_sampler_partitions_and_epochs():
    from b200dist.utils import DistributedSampler

    data = list(range(10))
    parts = [list(DistributedSampler(data, num_replicas=4, rank=r, shuffle=False)) for r in range(4)]
    assert all(len(p) == 3 for p in parts)
    assert sorted(sum(parts, []))[:10] == sorted(data + [0, 1])[:10]
    parts = [list(DistributedSampler(data, num_replicas=4, rank=r, shuffle=False, drop_last=True)) for r in range(4)]
    assert all(len(p) == 2 for p in parts) and len(set(sum(parts, []))) == 8
    s = DistributedSampler(data, num_replicas=2, rank=0, shuffle=True, seed=1)
    a = list(s); s.set_epoch(1); b = list(s); s.set_epoch(0)
    assert a != b and list(s) == a
    ref = torch.utils.data.DistributedSampler(data, num_replicas=2, rank=1, shuffle=True, seed=1)
    mine = DistributedSampler(data, num_replicas=2, rank=1, shuffle=True, seed=1)
    assert list(ref) == list(mine)


def test_checkpoint_cross_loads_with_reference(tmp_path):
    """A file we write loads with the reference's loader and vice versa."""
    if not (REF / "utils" / "checkpointing.py").exists():
        pytest.skip("reference tree not mounted")
    import importlib.util

    spec = importlib.util.spec_from_file_location("ref_ckpt", REF / "utils" / "checkpointing.py")
    ref_ckpt = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_ckpt)
    from b200dist.utils import load_checkpoint, save_checkpoint

    torch.manual_seed(0)
    m1 = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 2))
    o1 = torch.optim.SGD(m1.parameters(), lr=0.1, momentum=0.9)
    m1(torch.randn(3, 4)).sum().backward(); o1.step()
    p = str(tmp_path / "ours.pt")
    save_checkpoint(m1, o1, 7, p)
    blob = torch.load(p, weights_only=False)
    assert set(blob) == {"model_state_dict", "optimizer_state_dict", "epoch"}
    m2 = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 2))
    o2 = torch.optim.SGD(m2.parameters(), lr=0.1, momentum=0.9)
    assert ref_ckpt.load_checkpoint(m2, o2, p) == 7
    assert all(torch.equal(a, b) for a, b in zip(m1.state_dict().values(), m2.state_dict().values()))
    p2 = str(tmp_path / "theirs.pt")
    ref_ckpt.save_checkpoint(m1, o1, 9, p2)
    m3 = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 2))
    assert load_checkpoint(m3, None, p2) == 9
    assert all(torch.equal(a, b) for a, b in zip(m1.state_dict().values(), m3.state_dict().values()))
    assert load_checkpoint(m3, None, str(tmp_path / "missing.pt")) == 0


def test_snapshot_and_sharded_roundtrip(tmp_path):
    from b200dist.utils import (consolidate_sharded_checkpoint, load_snapshot, save_sharded_checkpoint, save_snapshot)

    m = nn.Linear(3, 3)
    save_snapshot(m, 4, str(tmp_path / "snapshot.pt"))
    snap = torch.load(tmp_path / "snapshot.pt", weights_only=False)
    assert set(snap) == {"MODEL_STATE", "EPOCHS_RUN"}
    m2 = nn.Linear(3, 3)
    assert load_snapshot(m2, str(tmp_path / "snapshot.pt")) == 4 and torch.equal(m.weight, m2.weight)
    save_sharded_checkpoint(m, None, 2, str(tmp_path / "sh"))
    consolidate_sharded_checkpoint(str(tmp_path / "sh"), str(tmp_path / "full.pt"))
    full = torch.load(tmp_path / "full.pt", weights_only=False)
    assert torch.equal(full["model_state_dict"]["weight"], m.weight) and full["epoch"] == 2


def test_llama_state_dict_compat_with_reference():
    if not (REF / "fsdp_tp" / "llama2_model.py").exists():
        pytest.skip("reference tree not mounted")
    sys.path.insert(0, str(REF / "fsdp_tp"))
    try:
        import llama2_model as ref
    finally:
        sys.path.pop(0)
    from b200dist.models.llama2 import ModelArgs, Transformer, ffn_hidden_dim

    a = dict(dim=64, n_layers=2, n_heads=4, n_kv_heads=2, vocab_size=128, max_seq_len=32, multiple_of=32)
    torch.manual_seed(0)
    r = ref.Transformer.from_model_args(ref.ModelArgs(**a))
    m = Transformer(ModelArgs(**a))
    assert list(r.state_dict().keys()) == list(m.state_dict().keys())
    assert all(r.state_dict()[k].shape == m.state_dict()[k].shape and r.state_dict()[k].dtype == m.state_dict()[k].dtype
               for k in r.state_dict())
    m.load_state_dict(r.state_dict())
    tok = torch.randint(0, 128, (2, 16))
    yr, ym = r(tok), m(tok)
    assert ym.dtype == torch.float32 and torch.allclose(yr, ym, atol=1e-5)
    yr.sum().backward(); ym.sum().backward()
    for (n, p), q in zip(r.named_parameters(), m.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-4, rtol=1e-3), n
    assert ffn_hidden_dim(4096, 256, None) == 11008 and ffn_hidden_dim(5120, 256, None) == 13824
    assert m.to(torch.bfloat16).freqs_cis.dtype == torch.complex64


def _order_worker(rank, world):
    from b200dist.utils.debug import CollectiveOrderChecker, CollectiveOrderError, assert_replicas_equal

    c = CollectiveOrderChecker()
    for i in range(5):
        c.record(0, "all_reduce", 1024)
        c.record(2, "rs_adamw_from", 4096 + i)
    c.verify()                                   # identical sequences on both ranks: passes
    if rank == 1:
        c.record(2, "all_gather_from", 64)       # rank 1 issues one extra collective on channel 2
    try:
        c.verify()
    except CollectiveOrderError as e:
        assert "channel 2" in str(e) and "rank 1" in str(e)
    else:
        raise AssertionError("diverging collective order was not detected")
    t = torch.arange(16, dtype=torch.float32)
    assert_replicas_equal(t, name="same")
    try:
        assert_replicas_equal(t + (rank == 1) * 1e-3, name="diverged")
    except AssertionError as e:
        assert "diverged" in str(e)
    else:
        raise AssertionError("replica divergence was not detected")


def test_collective_order_hash_and_replica_check_gloo_world2():
    run_distributed(_order_worker, 2)


def test_stream_hazard_tracker():
    import pytest

    from b200dist.utils.debug import StreamHazardError, StreamHazardTracker

    t = StreamHazardTracker()
    t.write("stage0", "main")                    # backward writes gradients into the staging buffer
    t.record("grads_ready", "main")
    t.wait("grads_ready", "comm")
    t.read("stage0", "comm")                     # reduce-scatter reads it after the event: fine
    with pytest.raises(StreamHazardError, match="write-after-read"):
        t.write("stage0", "main")                # next backward overwrites while comm may still be reading
    t2 = StreamHazardTracker()
    t2.write("flat", "prefetch")
    with pytest.raises(StreamHazardError, match="read-after-write"):
        t2.read("flat", "main")                  # compute consumes the gathered weights without waiting for the gather
    t2.record("gathered", "prefetch"); t2.wait("gathered", "main")
    t2.read("flat", "main")
    t2.record("consumed", "main"); t2.wait("consumed", "prefetch")
    t2.write("flat", "prefetch")                 # properly ordered reuse
    with pytest.raises(StreamHazardError, match="never recorded"):
        t2.wait("missing", "main")


def test_sync_debug_toggle():
    from b200dist.utils import debug

    assert not debug.sync_debug_enabled()
    with debug.sync_debug():
        assert debug.sync_debug_enabled()
    assert not debug.sync_debug_enabled()


def test_fault_injection_and_resume_from_checkpoint(tmp_path, monkeypatch):
    """Restart-from-checkpoint drill: a fault injected before step 2 kills the run, a fresh Trainer resumes from the
    last checkpoint and ends with exactly the weights of an uninterrupted run."""
    import pytest

    from b200dist.engine import InjectedFault, Trainer
    from b200dist.utils import load_checkpoint, save_checkpoint

    def make():
        torch.manual_seed(0)
        m = torch.nn.Linear(8, 4)
        return m, torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9)

    data = [(torch.randn(5, 8, generator=torch.Generator().manual_seed(i)), torch.randint(0, 4, (5,), generator=torch.Generator().manual_seed(i)))
            for i in range(4)]
    loss = torch.nn.functional.cross_entropy
    m_ref, o_ref = make()
    t = Trainer(m_ref, o_ref, loss_fn=loss)
    for x, y in data:
        t.train_step(x, y)
    ck = str(tmp_path / "ck.pt")
    monkeypatch.setenv("B200DIST_INJECT_FAULT", "0:2")
    m, o = make()
    t = Trainer(m, o, loss_fn=loss)
    with pytest.raises(InjectedFault):
        for i, (x, y) in enumerate(data):
            t.train_step(x, y)
            save_checkpoint(m, o, i + 1, ck)
    monkeypatch.delenv("B200DIST_INJECT_FAULT")
    m2, o2 = make()
    done = load_checkpoint(m2, o2, ck)
    assert done == 2
    t2 = Trainer(m2, o2, loss_fn=loss)
    for x, y in data[done:]:
        t2.train_step(x, y)
    assert torch.allclose(m2.weight, m_ref.weight) and torch.allclose(m2.bias, m_ref.bias)


def test_public_header_declares_every_exported_symbol():
    """include/b200dist.h is the C ABI contract: every b200_* symbol the library exports must be declared there."""
    import re
    import subprocess
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    so = root / "b200dist" / "_native" / "libb200dist.so"
    if not so.exists():
        import pytest

        pytest.skip("native library not built")
    out = subprocess.run(["nm", "-D", "--defined-only", str(so)], capture_output=True, text=True, check=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T b200_" in l}
    declared = set(re.findall(r"\bb200_[a-z0-9_]+", (root / "include" / "b200dist.h").read_text()))
    assert exported and not (exported - declared), sorted(exported - declared)


def test_capturable_adamw_matches_torch_and_tracks_lr():
    """capturable=True keeps (lr, bias corrections) in a per-group tensor refreshed by prepare_step(): same numbers as
    torch.optim.AdamW, including after a learning-rate change (the values a replayed CUDA graph would read)."""
    from b200dist.ops import FusedAdamW

    torch.manual_seed(0)
    a = torch.nn.Parameter(torch.randn(6, 5)); b = torch.nn.Parameter(a.detach().clone())
    oa = FusedAdamW([a], lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1, capturable=True)
    ob = torch.optim.AdamW([b], lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1)
    for i in range(5):
        if i == 3:
            oa.param_groups[0]["lr"] = ob.param_groups[0]["lr"] = 3e-3
        g = torch.randn(6, 5)
        a.grad, b.grad = g.clone(), g.clone()
        oa.step(); ob.step()
    assert torch.allclose(a, b, atol=1e-6)
    sc = oa.param_groups[0]["_scalars"]
    assert oa.param_groups[0]["_t"] == 5 and abs(float(sc[0]) - 3e-3) < 1e-9
    assert abs(float(sc[1]) - (1 - 0.9 ** 5)) < 1e-6 and abs(float(sc[2]) - (1 - 0.95 ** 5) ** 0.5) < 1e-6
    # a trainer asked for a graph on a CPU model simply runs eagerly
    from b200dist.engine import Trainer

    m = torch.nn.Linear(4, 3)
    t = Trainer(m, FusedAdamW(m.parameters(), capturable=True), loss_fn=torch.nn.functional.cross_entropy, cuda_graph=True)
    t.train_step(torch.randn(2, 4), torch.tensor([0, 2]))
    assert t._graph is None and t.steps_done == 1


def _single_box_worker(rank, world):
    import os

    import torch

    from b200dist.comm.collectives import _agree_single_box

    ok, why = _agree_single_box(None, torch.device("cpu"), world)
    assert ok and why == "", (ok, why)
    if rank == 1:
        os.environ["B200DIST_FAKE_HOST"] = "another-node"
    ok, why = _agree_single_box(None, torch.device("cpu"), world)   # every rank must reach the SAME verdict
    assert not ok and "different hosts" in why, (ok, why)


def test_communicator_refuses_peer_memory_across_hosts():
    """a group whose ranks are not on one box must not exchange CUDA-IPC handles; all ranks agree to fall back."""
    from tests.dist_utils import run_distributed

    run_distributed(_single_box_worker, 2)


@pytest.mark.skipif(not REF.exists(), reason="reference tree not mounted")
def test_reference_arm_provisions_itself():
    """baseline/install_ref.py creates baseline/_ref (pip --target + verbatim recipe trees) on demand and is idempotent."""
    import filecmp

    from baseline import install_ref as ir

    root = ir.ensure_reference()
    assert root is not None, ir.why_unavailable()
    assert (root / "utils" / "__init__.py").exists()
    for rel in ("fsdp_tp/llama2_model.py", "scripts/main.py", "tests/torch_comm_bench.py"):
        assert filecmp.cmp(REF / rel, ir.reference_path(*rel.split("/")), shallow=False), rel
    assert ir.ensure_reference() == root


def test_attention_cta_order_covers_every_tile_once():
    """Host restatement of ``lpt_decode`` (csrc/attn_common.cuh): the grouped heavy-first reading of the linear CTA index must
    hit every (tile, head, batch) exactly once, also when heads x batch is not a multiple of the group size (40 heads in
    groups of 16), and inside a group all heads' tiles of one weight class come before any lighter tile."""
    def decode(bx, by, bz, nx, ny, nz, group):
        per = ny * nz
        lin = bx + nx * (by + ny * bz)
        full = nx * group
        g = lin // full
        base = g * group
        gsz = min(group, per - base)
        within = lin - g * full
        rank = within // gsz
        hb = base + (within - rank * gsz)
        return rank, hb % ny, hb // ny, lin

    for nx, ny, nz, group in ((16, 32, 1, 16), (32, 40, 1, 8), (16, 40, 2, 16), (3, 10, 1, 16), (1, 4, 3, 8), (32, 32, 2, 8)):
        seen, order = set(), []
        for bz in range(nz):
            for by in range(ny):
                for bx in range(nx):
                    rank, h, b, lin = decode(bx, by, bz, nx, ny, nz, group)
                    assert 0 <= rank < nx and 0 <= h < ny and 0 <= b < nz
                    seen.add((rank, h, b))
                    order.append((lin, rank, b * ny + h))
        assert len(seen) == nx * ny * nz
        order.sort()
        # dispatch order: groups ascend; inside a group the rank (heavy -> light) never decreases
        last = {}
        for _, rank, hb in order:
            g = hb // group
            assert rank >= last.get(g, 0)
            last[g] = rank
