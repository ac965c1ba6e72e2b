"""Single-GPU numerical oracles: every sm_100a kernel vs a plain fp32 PyTorch expression of the same op
(SURVEY.md §4 item 2).  The checks themselves live in tools/gpu_check*.py so the same code serves bring-up runs."""
import pytest
import torch

pytestmark = pytest.mark.gpu

GROUPS = ["gemm_k1", "gemm_k2", "gemm_k3", "simt", "norms", "elementwise", "attn_fwd", "attn_bwd", "conv", "models", "optim_overlap", "fused_block", "attn_dropout"]


@pytest.mark.parametrize("group", GROUPS)
def test_kernel_group(group):
    from b200dist.ops import _lib
    from tools import gpu_check

    assert _lib.available(), "native library missing on a GPU box: run `python -m b200dist._build`"
    before = _lib.launch_count
    results = []
    gpu_check.GROUPS[group](results)
    torch.cuda.synchronize()
    failed = [r for r in results if not r["ok"]]
    assert not failed, f"{len(failed)} of {len(results)} checks failed: {failed[:5]}"
    assert _lib.launch_count > before, "no native kernel was launched (silent fallback?)"


def test_ops_fail_loudly_without_fallback():
    """CUDA tensors never take the eager PyTorch path: unsupported inputs raise instead."""
    from b200dist import ops

    q = torch.randn(1, 8, 2, 256, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(NotImplementedError):
        ops.attention(q, q, q)


def test_smoke_entry():
    import __graft_entry__ as g

    g.smoke()


def test_conv_implicit_gemm_fprop_dgrad_wgrad():
    """Implicit-GEMM convolution (4-D tensor-map operands) vs F.conv2d in fp32: forward, input gradient (flipped filter)
    and weight gradient (pixel-box split-K) over ResNet / UNet shapes incl. odd extents; tools/conv_igemm_check.py."""
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, str(root / "tools" / "conv_igemm_check.py")], capture_output=True, text=True,
                       timeout=300, cwd=str(root))
    assert r.returncode == 0 and "failures 0" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.skipif(__import__("os").environ.get("B200DIST_TEST_CUDA_GRAPH", "1") == "0",
                    reason="CUDA-graph training step check switched off (B200DIST_TEST_CUDA_GRAPH=0)")
def test_cuda_graph_training_step_matches_eager():
    """Runs in its own process (``tools/cuda_graph_check.py``): same loss sequence as the eager step, graph captured."""
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, str(root / "tools" / "cuda_graph_check.py")], capture_output=True, text=True,
                       timeout=300, cwd=str(root))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
