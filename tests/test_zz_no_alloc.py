"""b200_rmsnorm_bwd_noalloc, exercised straight through the C ABI.

The entry point exists for callers that cannot give the library a scratch buffer: a host program with no
device allocator of its own, a job whose HBM is already fully handed out to pools, or a loop where
cudaMalloc / cudaFree are forbidden because they stall the whole device.  Its scratch is a fixed array
inside the library, so this test checks two things a user of it relies on: (1) it produces what the regular
entry point (caller-provided scratch) produces, within the tolerance the other kernel tests use, and
(2) that still holds on a second call in a row with a different shape, i.e. whatever the first call left in
the shared array does not leak into the next result.  It also checks the documented size limit is enforced
with a clear error.  The file name keeps it last in pytest's collection order.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

I, F = C.c_int, C.c_float


def _ptr(t):
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def _run_both(lib, rows, D, dtype, seed):
    dev = "cuda"
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(rows, D, generator=g).to(dev).to(dtype)
    dy = torch.randn(rows, D, generator=g).to(dev).to(dtype)
    w = (1 + 0.1 * torch.randn(D, generator=g)).to(dev).to(dtype)
    rstd = torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5).contiguous()
    code = 0 if dtype == torch.bfloat16 else 1
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # existing entry point: the caller owns the scratch
    parts = lib.b200_norm_bwd_partials(I(rows))
    scratch = torch.empty(parts * D, dtype=torch.float32, device=dev)
    dx_ref, dw_ref = torch.empty_like(x), torch.empty_like(w)
    rc = lib.b200_rmsnorm_bwd(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dx_ref), _ptr(dw_ref), _ptr(scratch), I(rows),
                              I(D), I(0), I(0), I(code), stream)
    assert rc == 0, lib.b200_last_error()
    # new entry point: nothing is handed over, nothing is allocated
    dx, dw = torch.full_like(x, float("nan")), torch.full_like(w, float("nan"))
    rc = lib.b200_rmsnorm_bwd_noalloc(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dx), _ptr(dw), I(rows), I(D), I(0), I(0),
                                      I(code), stream)
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    return dx, dx_ref, dw, dw_ref


def test_noalloc_rmsnorm_bwd_matches_regular_entry_point_twice():
    from b200dist.ops import _lib

    lib = _lib.load()
    lib.b200_last_error.restype = C.c_char_p
    # two calls in a row on the same device: a wide one, then a narrower one with fewer rows, so the second
    # call runs over scratch the first one has already dirtied
    for rows, D, dtype, tol in ((777, 4096, torch.bfloat16, 3e-2), (50, 256, torch.float32, 1e-4)):
        dx, dx_ref, dw, dw_ref = _run_both(lib, rows, D, dtype, seed=rows)
        assert torch.isfinite(dx.float()).all() and torch.isfinite(dw.float()).all()
        assert torch.allclose(dx.float(), dx_ref.float(), atol=tol, rtol=tol)
        # dw sums `rows` terms; the two entry points split the rows over a different number of partials
        assert torch.allclose(dw.float(), dw_ref.float(), atol=tol * rows ** 0.5, rtol=tol)


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    from b200dist.ops import _lib

    lib = _lib.load()
    lib.b200_last_error.restype = C.c_char_p
    D = 8192 + 8
    x = torch.zeros(4, D, device="cuda", dtype=torch.bfloat16)
    rstd = torch.ones(4, device="cuda")
    rc = lib.b200_rmsnorm_bwd_noalloc(_ptr(x), _ptr(x), _ptr(x[0]), _ptr(rstd), _ptr(x), _ptr(x[0]), I(4), I(D), I(0), I(0),
                                      I(0), C.c_void_p(0))
    assert rc == 4  # B200_ERR_LIMIT
    assert b"exceeds the built-in scratch limit" in lib.b200_last_error()
