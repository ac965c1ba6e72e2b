"""Bring-up harness: runs kernel groups in separate subprocesses (a fault or hang in
one group cannot take the others down), compares against fp32 PyTorch references and
times with CUDA events.  Usage on a B200:

    python tools/gpu_check.py            # all groups
    python tools/gpu_check.py gemm_k1    # one group, in-process

Writes gpurun_out/check_<group>.log and a summary gpurun_out/check_summary.json.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
OUT = ROOT / "gpurun_out"


def _time(fn, iters=10, warmup=3, flush=None):
    import torch

    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def _gemm_raw(A, B, D, M, N, K, a_mn, b_mn, config, bias=None, residual=None, alpha=1.0, act=0,
              accumulate=0, max_ctas=0):
    import ctypes as C
    import torch
    from b200dist.ops import _lib as L

    lda = A.stride(0)
    ldb = B.stride(0)
    L.call("gemm_bf16", L.ptr(A), L.ptr(B), L.ptr(D), C.c_int(M), C.c_int(N), C.c_int(K),
           C.c_longlong(lda), C.c_longlong(ldb), C.c_longlong(D.stride(0)), C.c_int(a_mn), C.c_int(b_mn),
           C.c_int(1 if D.dtype == torch.float32 else 0), L.ptr(bias), L.ptr(residual),
           C.c_longlong(residual.stride(0) if residual is not None else 0), C.c_float(alpha),
           C.c_int(act), C.c_int(accumulate), C.c_int(config), C.c_int(max_ctas), L.stream_ptr())


def _check(name, got, ref, atol, rtol, results):
    import torch

    got = got.float()
    ref = ref.float()
    err = (got - ref).abs()
    tol = atol + rtol * ref.abs()
    bad = int((err > tol).sum().item())
    ok = bad == 0 and bool(torch.isfinite(got).all().item())
    results.append({"name": name, "ok": ok, "max_err": float(err.max().item()), "bad": bad,
                    "ref_absmax": float(ref.abs().max().item())})
    print(f"[{'PASS' if ok else 'FAIL'}] {name}: max_err={err.max().item():.4g} bad={bad}/{err.numel()} "
          f"ref_absmax={ref.abs().max().item():.4g}", flush=True)
    return ok


def group_gemm(config: int, results):
    import torch

    torch.manual_seed(0)
    dev = "cuda"
    shapes = [(256, 256, 128), (128, 128, 64), (384, 512, 320), (1000, 520, 264), (4096, 4096, 4096)]
    for (M, N, K) in shapes:
        for a_mn, b_mn in [(0, 0), (0, 1), (1, 1), (1, 0)]:
            A = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
            B = (torch.randn(N, K, device=dev) * 0.5).bfloat16()
            ref = A.float() @ B.float().t()
            As = A.t().contiguous() if a_mn else A
            Bs = B.t().contiguous() if b_mn else B
            for out_dt in (torch.bfloat16, torch.float32):
                D = torch.full((M, N), float("nan"), device=dev, dtype=out_dt)
                _gemm_raw(As, Bs, D, M, N, K, a_mn, b_mn, config)
                torch.cuda.synchronize()
                tol = 2e-2 if out_dt == torch.bfloat16 else 2e-3
                _check(f"gemm cfg{config} {M}x{N}x{K} a_mn={a_mn} b_mn={b_mn} out={str(out_dt)[6:]}", D, ref,
                       atol=tol * max(1.0, K ** 0.5 * 0.25), rtol=1e-2, results=results)
    # epilogue: bias + gelu + residual + accumulate
    M, N, K = 512, 768, 256
    A = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
    B = (torch.randn(N, K, device=dev) * 0.5).bfloat16()
    bias = torch.randn(N, device=dev).bfloat16()
    res = torch.randn(M, N, device=dev).bfloat16()
    D0 = torch.randn(M, N, device=dev)
    D = D0.clone()
    _gemm_raw(A, B, D, M, N, K, 0, 0, config, bias=bias, residual=res, alpha=0.5, act=1, accumulate=1)
    ref = torch.nn.functional.gelu(0.5 * (A.float() @ B.float().t()) + bias.float()) + res.float() + D0
    _check(f"gemm cfg{config} epilogue bias+gelu+res+acc", D, ref, atol=5e-2, rtol=1e-2, results=results)
    # timing on Llama-7B shapes
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    for (M, N, K) in [(8192, 4096, 4096), (8192, 11008, 4096), (8192, 4096, 11008), (4096, 32000, 4096)]:
        A = torch.randn(M, K, device=dev).bfloat16()
        B = torch.randn(N, K, device=dev).bfloat16()
        D = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        ms = _time(lambda: _gemm_raw(A, B, D, M, N, K, 0, 0, config), flush=flush)
        ms_ref = _time(lambda: torch.matmul(A, B.t(), out=D), flush=flush)
        tf = 2.0 * M * N * K / ms / 1e9
        tfr = 2.0 * M * N * K / ms_ref / 1e9
        print(f"[TIME] gemm cfg{config} {M}x{N}x{K}: {ms:.3f} ms {tf:.0f} TFLOP/s | cuBLAS {ms_ref:.3f} ms {tfr:.0f} TFLOP/s "
              f"| ratio {tf / tfr:.2f}", flush=True)
        results.append({"name": f"time gemm cfg{config} {M}x{N}x{K}", "ok": True, "ms": ms, "tflops": tf,
                        "cublas_ms": ms_ref, "cublas_tflops": tfr})
    # backward layouts timing (dgrad: K-major x MN-major; wgrad: MN x MN)
    M, N, K = 8192, 4096, 4096
    A = torch.randn(M, K, device=dev).bfloat16()
    Bt = torch.randn(K, N, device=dev).bfloat16()
    D = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    ms = _time(lambda: _gemm_raw(A, Bt, D, M, N, K, 0, 1, config), flush=flush)
    print(f"[TIME] gemm cfg{config} dgrad-layout {M}x{N}x{K}: {ms:.3f} ms {2.0*M*N*K/ms/1e9:.0f} TFLOP/s", flush=True)
    At = torch.randn(K, M, device=dev).bfloat16()
    Df = torch.empty(M, N, device=dev, dtype=torch.float32)
    ms = _time(lambda: _gemm_raw(At, Bt, Df, M, N, K, 1, 1, config), flush=flush)
    print(f"[TIME] gemm cfg{config} wgrad-layout fp32-out {M}x{N}x{K}: {ms:.3f} ms {2.0*M*N*K/ms/1e9:.0f} TFLOP/s", flush=True)


def group_simt(results):
    import torch

    dev = "cuda"
    # split-K path: skinny fp32 outputs with a long reduction (conv / small-layer weight gradients)
    torch.manual_seed(1)
    for (M, N, K, amn, bmn) in [(64, 576, 50048, 1, 1), (128, 128, 16384, 0, 0), (256, 1152, 25088, 1, 1)]:
        A = (torch.randn(M, K, device=dev) * 0.1).bfloat16()
        B = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
        ref = A.float() @ B.float().t()
        As = A.t().contiguous() if amn else A
        Bs = B.t().contiguous() if bmn else B
        D = torch.full((M, N), float("nan"), device=dev, dtype=torch.float32)
        _gemm_raw(As, Bs, D, M, N, K, amn, bmn, 0)
        _check(f"gemm split-K {M}x{N}x{K} a_mn={amn}", D, ref, 5e-2, 1e-2, results)
        D2 = torch.ones((M, N), device=dev, dtype=torch.float32)
        _gemm_raw(As, Bs, D2, M, N, K, amn, bmn, 0, accumulate=1)
        _check(f"gemm split-K accumulate {M}x{N}x{K}", D2, ref + 1, 5e-2, 1e-2, results)
        ms = _time(lambda: _gemm_raw(As, Bs, D, M, N, K, amn, bmn, 0))
        print(f"[TIME] gemm split-K {M}x{N}x{K}: {ms*1e3:.1f} us {2.0*M*N*K/ms/1e9:.0f} TFLOP/s", flush=True)
    for (M, N, K) in [(32, 64, 10), (7, 2, 64), (50, 16, 20)]:
        A = torch.randn(M, K, device=dev).bfloat16()
        B = torch.randn(N, K, device=dev).bfloat16()
        D = torch.empty(M, N, device=dev, dtype=torch.float32)
        _gemm_raw(A, B, D, M, N, K, 0, 0, 9)
        _check(f"gemm simt {M}x{N}x{K}", D, A.float() @ B.float().t(), 1e-2, 1e-2, results)


GROUPS = {
    "gemm_k1": lambda r: group_gemm(1, r),
    "gemm_k2": lambda r: group_gemm(2, r),
    "gemm_k3": lambda r: group_gemm(3, r),
    "simt": group_simt,
}
try:
    from tools import gpu_check_more as _more

    GROUPS.update(_more.GROUPS)
except ImportError:
    pass


def run_group(name: str) -> int:
    import torch

    results = []
    t0 = time.time()
    try:
        GROUPS[name](results)
        torch.cuda.synchronize()
        err = None
    except Exception as e:  # noqa: BLE001
        import traceback

        traceback.print_exc()
        err = repr(e)
    OUT.mkdir(exist_ok=True)
    nfail = sum(1 for r in results if not r["ok"])
    (OUT / f"check_{name}.json").write_text(json.dumps(
        {"group": name, "error": err, "n": len(results), "n_fail": nfail, "secs": time.time() - t0,
         "results": results}, indent=1))
    print(f"== group {name}: {len(results)} checks, {nfail} failed, error={err}", flush=True)
    return 1 if (err or nfail) else 0


def main():
    OUT.mkdir(exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] != "--all" and not sys.argv[1].startswith("--only="):
        sys.exit(run_group(sys.argv[1]))
    names = list(GROUPS)
    if len(sys.argv) > 1 and sys.argv[1].startswith("--only="):
        names = sys.argv[1][7:].split(",")
    extra = []
    summary = {}
    for n in names + extra:
        log = OUT / f"check_{n}.log"
        t0 = time.time()
        with open(log, "w") as f:
            try:
                rc = subprocess.run([sys.executable, __file__, n], stdout=f, stderr=subprocess.STDOUT,
                                    timeout=int(os.environ.get("CHECK_TIMEOUT", "240"))).returncode
            except subprocess.TimeoutExpired:
                rc = -999
        summary[n] = {"rc": rc, "secs": round(time.time() - t0, 1)}
        tail = log.read_text().splitlines()
        fails = [l for l in tail if "FAIL" in l or "Error" in l or "error" in l][:12]
        times = [l for l in tail if l.startswith("[TIME]")]
        print(f"### {n}: rc={rc} ({summary[n]['secs']}s)")
        for l in fails + times + tail[-3:]:
            print("   ", l)
    (OUT / "check_summary.json").write_text(json.dumps(summary, indent=1))
    sys.exit(0)


if __name__ == "__main__":
    main()
