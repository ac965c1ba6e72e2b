"""Ulysses / ring attention / halo-exchange convolution vs their unsharded single-process results (gloo, CPU)."""
import torch
import torch.nn.functional as F

from tests.dist_utils import run_distributed


def _attn_inputs(B=2, S=16, H=4, Hkv=4, hd=8):
    g = torch.Generator().manual_seed(0)
    return (torch.randn(B, S, H, hd, generator=g), torch.randn(B, S, Hkv, hd, generator=g),
            torch.randn(B, S, Hkv, hd, generator=g), torch.randn(B, S, H, hd, generator=g))


def _ctx_worker(rank, world, kind, causal, Hkv):
    from b200dist import ops
    from b200dist.comm import get_communicator
    from b200dist.parallel import ring_attention, ulysses_attention

    comm = get_communicator(None)
    q, k, v, do = _attn_inputs(Hkv=Hkv)
    qr, kr, vr = [t.clone().requires_grad_(True) for t in (q, k, v)]
    ref = ops.attention(qr, kr, vr, causal=causal)
    ref.backward(do)
    sl = slice(rank * 16 // world, (rank + 1) * 16 // world)
    ql, kl, vl = [t[:, sl].clone().requires_grad_(True) for t in (q, k, v)]
    fn = ring_attention if kind == "ring" else ulysses_attention
    out = fn(ql, kl, vl, comm, causal=causal)
    assert torch.allclose(out, ref[:, sl], atol=1e-5), (out - ref[:, sl]).abs().max()
    out.backward(do[:, sl])
    for a, b, n in ((ql.grad, qr.grad[:, sl], "dq"), (kl.grad, kr.grad[:, sl], "dk"), (vl.grad, vr.grad[:, sl], "dv")):
        assert torch.allclose(a, b, atol=1e-5), (n, (a - b).abs().max())


def test_ring_attention_causal():
    run_distributed(_ctx_worker, 4, args=("ring", True, 4))


def test_ring_attention_noncausal_gqa():
    run_distributed(_ctx_worker, 2, args=("ring", False, 2))


def test_ulysses_attention_causal():
    run_distributed(_ctx_worker, 2, args=("ulysses", True, 4))


def test_ulysses_attention_noncausal():
    run_distributed(_ctx_worker, 4, args=("ulysses", False, 4))


def _halo_worker(rank, world):
    from b200dist.comm import get_communicator
    from b200dist.parallel import halo_conv2d

    comm = get_communicator(None)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 3, 12, 7, generator=g)
    w = torch.randn(5, 3, 3, 3, generator=g)
    b = torch.randn(5, generator=g)
    dy = torch.randn(2, 5, 12, 7, generator=g)
    xr, wr = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    ref = F.conv2d(xr, wr, b, padding=1)
    ref.backward(dy)
    rows = slice(rank * 12 // world, (rank + 1) * 12 // world)
    xl, wl = x[:, :, rows].clone().requires_grad_(True), w.clone().requires_grad_(True)
    out = halo_conv2d(xl, wl, b, comm, padding=1)
    assert torch.allclose(out, ref[:, :, rows], atol=1e-5)
    out.backward(dy[:, :, rows])
    assert torch.allclose(xl.grad, xr.grad[:, :, rows], atol=1e-5), (xl.grad - xr.grad[:, :, rows]).abs().max()
    import torch.distributed as dist

    gw = wl.grad.clone()
    dist.all_reduce(gw)  # weight gradients are summed over the spatial shards
    assert torch.allclose(gw, wr.grad, atol=1e-4)
    # the naive split (no halo) is wrong exactly at the cut
    naive = F.conv2d(x[:, :, rows], w, b, padding=1)
    if world > 1:
        assert not torch.allclose(naive, ref[:, :, rows], atol=1e-3)


def test_halo_conv_matches_full_conv():
    run_distributed(_halo_worker, 3)


def _zigzag_worker(rank, world, Hkv):
    from b200dist import ops
    from b200dist.comm import get_communicator
    from b200dist.parallel.ring_attention import (_ZigzagRingAttn, zigzag_positions, zigzag_ring_attention, zigzag_shard,
                                                  zigzag_unshard)

    comm = get_communicator(None)
    q, k, v, do = _attn_inputs(Hkv=Hkv)
    qr, kr, vr = [t.clone().requires_grad_(True) for t in (q, k, v)]
    ref = ops.attention(qr, kr, vr, causal=True)
    ref.backward(do)
    ql, kl, vl = [zigzag_shard(t, world, rank).requires_grad_(True) for t in (q, k, v)]
    out = zigzag_ring_attention(ql, kl, vl, comm)
    assert torch.allclose(out, zigzag_shard(ref.detach(), world, rank), atol=1e-5)
    out.backward(zigzag_shard(do, world, rank))
    for a, b, n in ((ql.grad, qr.grad, "dq"), (kl.grad, kr.grad, "dk"), (vl.grad, vr.grad, "dv")):
        assert torch.allclose(a, zigzag_shard(b, world, rank), atol=1e-5), (n, (a - zigzag_shard(b, world, rank)).abs().max())
    # layout helpers round-trip, and the work is balanced: every rank runs the same number of block pairs
    parts = [zigzag_shard(q, world, r) for r in range(world)]
    assert torch.equal(zigzag_unshard(parts), q)
    assert zigzag_positions(16, world, rank)[0][0] == rank * (16 // (2 * world))
    counts = [sum(len(_ZigzagRingAttn._pairs(world, r, t)[1]) for t in range(world)) for r in range(world)]
    assert len(set(counts)) == 1 and counts[0] == 2 * world + 1, counts


def test_zigzag_ring_attention_balanced_causal():
    run_distributed(_zigzag_worker, 2, args=(4,))


def test_zigzag_ring_attention_gqa_world4():
    run_distributed(_zigzag_worker, 4, args=(2,))
