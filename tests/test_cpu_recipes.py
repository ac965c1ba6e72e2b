"""Recipes end to end on CPU: BASELINE config 1 (ResNet-18 DDP, gloo, world 2, synthetic 224x224) and script smoke runs."""
import os
import subprocess
import sys
from pathlib import Path

import torch

from tests.dist_utils import free_port, run_distributed

ROOT = Path(__file__).resolve().parent.parent


def _resnet_ddp_worker(rank, world):
    from b200dist import ops
    from b200dist.models import build_resnet
    from b200dist.parallel import DDP

    torch.manual_seed(0)
    model = build_resnet("resnet18", num_classes=10)
    ddp = DDP(model)
    opt = torch.optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)
    g = torch.Generator().manual_seed(100 + rank)   # each rank trains on different synthetic images
    x, y = torch.rand(4, 3, 224, 224, generator=g), torch.randint(0, 10, (4,), generator=g)
    losses = []
    for _ in range(3):
        opt.zero_grad()
        loss = ops.cross_entropy(ddp(x), y)
        loss.backward()
        ddp.finish_backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0], losses
    import torch.distributed as dist

    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    other = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(other, flat)
    assert torch.equal(other[0], other[1]), "replicas diverged"           # bit-identical after N steps
    rm = model.bn1.running_mean.clone()
    others = [torch.empty_like(rm) for _ in range(world)]
    dist.all_gather(others, rm)
    # BN buffers follow rank 0 (broadcast_buffers=True): equal at the start of the last forward, local update after
    assert torch.isfinite(rm).all()


def test_config1_resnet18_ddp_gloo_world2():
    run_distributed(_resnet_ddp_worker, 2, timeout=600)


def _run(cmd, nproc=1, timeout=600):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), B200DIST_DISABLE_MPI4PY="1",
               OMP_NUM_THREADS="1")
    if nproc > 1:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
               "127.0.0.1", "--master-port", env["MASTER_PORT"]] + cmd
    else:
        cmd = [sys.executable] + cmd
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=str(ROOT), env=env)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    return r.stdout + r.stderr


def test_main_py_gloo_ddp(tmp_path):
    out = _run(["scripts/main.py", "--use_syn", "--arch", "resnet18", "--backend", "gloo", "--num_epochs", "2", "--steps_syn", "2",
                "--batch_size", "2", "--image_size", "64", "--num_classes", "10", "--logfile", str(tmp_path / "log.txt"),
                "--model_dir", str(tmp_path)], nproc=2)
    assert "images/s" in out and (tmp_path / "log.txt").read_text().count("Average epoch time") == 1


def _run_many(jobs, workers=4):
    """Run several (cmd, nproc) launches concurrently (each has its own rendezvous port); returns outputs in order."""
    from concurrent.futures import ThreadPoolExecutor

    with ThreadPoolExecutor(max_workers=workers) as ex:
        futs = [ex.submit(_run, cmd, nproc) for cmd, nproc in jobs]
        return [f.result() for f in futs]


def test_recipe_scripts_run_on_cpu(tmp_path):
    outs = _run_many([
        (["scripts/03_tensor_parallel_tp/01_device_mesh_basics.py"], 2),
        (["scripts/03_tensor_parallel_tp/02_basic_tensor_parallel.py"], 2),
        (["scripts/04_pipeline_parallel_pp/01_manual_model_split.py"], 2),
        (["scripts/04_pipeline_parallel_pp/03_pipeline_training.py", "--num-steps", "4", "--batch-size", "4", "--seq-len", "16",
          "--dim", "32", "--layers-per-stage", "1", "--vocab", "100", "--pp", "2"], 2),
        (["scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py", "--iters", "3", "--seq-len", "32", "--tp_size", "2"], 2),
        (["scripts/01_data_parallel_ddp/multinode_ddp_basic.py", "--total_epochs", "2", "--backend", "gloo", "--save_every", "1",
          "--snapshot", str(tmp_path / "snap.pt")], 2),
        (["scripts/07_domain_parallel/01_halo_conv.py"], 2),
    ])
    assert "✓" in outs[0]
    assert "iter 4" in outs[1]
    assert "pipeline output" in outs[2]
    assert "tokens/s" in outs[3]
    assert "tokens/s" in outs[4]
    assert (tmp_path / "snap.pt").exists()
    assert "with halo exchange" in outs[6]


def test_comm_bench_csv_columns(tmp_path):
    out = _run(["bench/comm_bench.py", "--backend", "gloo", "--tensor_sizes", "1024", "--warmup_runs", "1", "--benchmark_runs", "2",
                "--output_file", str(tmp_path / "r.log"), "--impl", "both"], nproc=2)
    text = (tmp_path / "r.log").read_text()
    assert "backend,operation,tensor_elements,tensor_bytes,tensor_mb,warmup_iterations" in text
    ops = {l.split(",")[1] for l in text.splitlines() if l and not l.startswith("#")}
    assert {"broadcast", "all_reduce", "all_gather", "reduce_scatter"} <= ops


def test_comm_tools_and_fsdp_tp_examples_run_on_cpu(tmp_path):
    outs = _run_many([
        (["tools/test_torchrun.py"], 2),
        (["tools/all_reduce_test.py", "--backend", "gloo", "--max-exp", "4", "--output", str(tmp_path / "ar.log")], 2),
        (["fsdp_tp/tensor_parallel_example.py", "--iters", "3"], 2),
        (["fsdp_tp/fsdp_tp_example.py", "--iters", "2", "--seq-len", "32", "--tp_size", "2"], 2),
        # import-compatible module names of the reference's fsdp_tp/ directory
        (["-c", "import sys; sys.path.insert(0, 'fsdp_tp'); from llama2_model import Transformer, ModelArgs, apply_rotary_emb; "
                "from log_utils import rank_log, get_logger, verify_min_gpu_count; print('imports ok')"], 1),
    ])
    assert "rank 1 has data 42" in outs[0]
    log = (tmp_path / "ar.log").read_text()
    assert log.count("all_reduce") == 2 and log.count("broadcast") == 2 and "elements=10000" in log
    assert "tensor-parallel example done" in outs[2]
    assert "tokens/s" in outs[3]
    assert "imports ok" in outs[4]
