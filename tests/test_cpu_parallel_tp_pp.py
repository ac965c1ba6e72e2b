"""TP/SP plans, hybrid TP x FSDP, and pipeline schedules on CPU (gloo)."""
import torch
import torch.nn as nn

from tests.dist_utils import run_distributed


def _toy_llama(seed=7):
    from b200dist.models.llama2 import ModelArgs, Transformer

    torch.manual_seed(seed)
    return Transformer(ModelArgs(dim=32, n_layers=2, n_heads=4, vocab_size=64, max_seq_len=32, multiple_of=16))


def _tp_worker(rank, world, sp):
    from b200dist import ops
    from b200dist.parallel import init_device_mesh
    from b200dist.parallel.tp import gather_tp_state_dict, llama_tp_plan

    ref = _toy_llama()
    model = _toy_llama()
    mesh = init_device_mesh("cpu", (world,), ("tp",))
    llama_tp_plan(model, mesh["tp"], sequence_parallel=sp)
    # sharded shapes
    assert model.layers[0].attention.wq.weight.shape == (32 // world, 32)
    assert model.layers[0].attention.wo.weight.shape == (32, 32 // world)
    assert model.tok_embeddings.weight.shape == (64 // world, 32)
    tok = torch.randint(0, 64, (2, 9), generator=torch.Generator().manual_seed(3))  # same on every TP rank
    x, y = tok[:, :-1], tok[:, 1:]
    out = model(x)
    out_ref = ref(x)
    assert torch.allclose(out, out_ref, atol=1e-4, rtol=1e-4), (out - out_ref).abs().max()
    ops.cross_entropy(out, y).backward()
    ops.cross_entropy(out_ref, y).backward()
    # compare a colwise, a rowwise, an embedding and a norm gradient against the unsharded model
    g = model.layers[1].feed_forward.w1.weight.grad
    gr = ref.layers[1].feed_forward.w1.weight.grad.chunk(world, 0)[rank]
    assert torch.allclose(g, gr, atol=1e-5, rtol=1e-3)
    g = model.layers[0].attention.wo.weight.grad
    gr = ref.layers[0].attention.wo.weight.grad.chunk(world, 1)[rank]
    assert torch.allclose(g, gr, atol=1e-5, rtol=1e-3)
    g = model.tok_embeddings.weight.grad
    gr = ref.tok_embeddings.weight.grad.chunk(world, 0)[rank]
    assert torch.allclose(g, gr, atol=1e-5, rtol=1e-3)
    g, gr = model.layers[0].ffn_norm.weight.grad, ref.layers[0].ffn_norm.weight.grad
    assert torch.allclose(g, gr, atol=1e-5, rtol=1e-3), (g - gr).abs().max()
    full = gather_tp_state_dict(model, mesh["tp"])
    for k, v in ref.state_dict().items():
        assert full[k].shape == v.shape, k
        if not v.is_complex():
            assert torch.allclose(full[k], v), k
    # and back: a full (reference-layout) state dict loads into the sharded module
    from b200dist.parallel.tp import load_tp_state_dict

    other = {k: (v + 1.0 if (v.is_floating_point() and not v.is_complex()) else v) for k, v in ref.state_dict().items()}
    load_tp_state_dict(model, mesh["tp"], other)
    again = gather_tp_state_dict(model, mesh["tp"])
    for k, v in other.items():
        if not v.is_complex():
            assert torch.allclose(again[k], v), k


def test_tp_sp_llama_matches_unsharded():
    run_distributed(_tp_worker, 2, args=(True,))


def test_tp_no_sp_llama_matches_unsharded():
    run_distributed(_tp_worker, 2, args=(False,))


def _basic_tp_worker(rank, world):
    """ToyMLP 16 -> 64 -> 16 with Colwise/Rowwise (reference 02_basic_tensor_parallel.py:46-71)."""
    from b200dist.parallel import init_device_mesh
    from b200dist.parallel.tp import ColwiseParallel, RowwiseParallel, parallelize_module

    class ToyMLP(nn.Module):
        def __init__(self):
            super().__init__()
            self.in_proj = nn.Linear(16, 64)
            self.relu = nn.ReLU()
            self.out_proj = nn.Linear(64, 16)

        def forward(self, x):
            return self.out_proj(self.relu(self.in_proj(x)))

    torch.manual_seed(0)
    ref = ToyMLP()
    torch.manual_seed(0)
    m = ToyMLP()
    mesh = init_device_mesh("cpu", (world,), ("tp",))
    parallelize_module(m, mesh, {"in_proj": ColwiseParallel(), "out_proj": RowwiseParallel()})
    x = torch.randn(5, 16, generator=torch.Generator().manual_seed(1), requires_grad=True)
    x2 = x.detach().clone().requires_grad_(True)
    y, yr = m(x), ref(x2)
    assert torch.allclose(y, yr, atol=1e-5)
    y.sum().backward(); yr.sum().backward()
    assert torch.allclose(x.grad, x2.grad, atol=1e-5)
    assert torch.allclose(m.out_proj.bias.grad, ref.out_proj.bias.grad, atol=1e-5)


def test_basic_tp_mlp():
    run_distributed(_basic_tp_worker, 2)


def _hybrid_worker(rank, world):
    """2-D mesh (dp=2, tp=2): TP first, then FSDP over dp — reference 01_fsdp_tp_hybrid.py:88-155."""
    from b200dist import ops
    from b200dist.parallel import FSDP, init_device_mesh
    from b200dist.parallel.tp import llama_tp_plan

    mesh = init_device_mesh("cpu", (2, 2), ("dp", "tp"))
    dp_rank, tp_rank = mesh.get_local_rank("dp"), mesh.get_local_rank("tp")
    assert mesh["tp"].size() == 2 and mesh["dp"].size() == 2
    model = _toy_llama()
    llama_tp_plan(model, mesh["tp"], sequence_parallel=True)
    fsdp = FSDP(model, process_group=mesh.get_group("dp"))
    opt = fsdp.build_optimizer("adamw", lr=1e-2)
    ref = _toy_llama()
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2)
    g = torch.Generator().manual_seed(5)
    for _ in range(2):
        tok = torch.randint(0, 64, (4, 9), generator=g)
        mine = tok.chunk(2)[dp_rank]  # data sharded by dp rank, replicated over tp
        opt.zero_grad()
        ops.cross_entropy(fsdp(mine[:, :-1]), mine[:, 1:]).backward()
        opt.step()
        ropt.zero_grad()
        ops.cross_entropy(ref(tok[:, :-1]), tok[:, 1:]).backward()
        ropt.step()
    sd = fsdp.full_state_dict(rank0_only=False)
    w = sd["layers.0.feed_forward.w1.weight"]
    wr = ref.layers[0].feed_forward.w1.weight.detach().chunk(2, 0)[tp_rank]
    assert torch.allclose(w, wr, atol=5e-5, rtol=1e-3), (w - wr).abs().max()


def test_hybrid_tp_fsdp_2d_mesh():
    run_distributed(_hybrid_worker, 4)


def test_schedule_actions_1f1b_and_gpipe():
    from b200dist.parallel.pp import bubble_fraction, schedule_actions

    S, M = 4, 8
    for stage in range(S):
        acts = schedule_actions("1f1b", stage, S, M)
        assert sorted(a for a in acts if a[0] == "F") == [("F", i) for i in range(M)]
        assert sorted(a for a in acts if a[0] == "B") == [("B", i) for i in range(M)]
        inflight = peak = 0
        for k, i in acts:
            inflight += 1 if k == "F" else -1
            peak = max(peak, inflight)
            assert inflight >= 0
        assert peak == min(M, S - stage), (stage, peak)  # 1F1B bounds live activations by the stage depth
        for i in range(M):  # a microbatch's backward never precedes its forward
            assert acts.index(("F", i)) < acts.index(("B", i))
    g = schedule_actions("gpipe", 1, S, M)
    assert g[:M] == [("F", i) for i in range(M)] and g[M:] == [("B", i) for i in range(M)]
    assert abs(bubble_fraction(4, 4) - 3 / 7) < 1e-12


def _pp_worker(rank, world, kind):
    from b200dist.parallel.pp import PipelineStage, Schedule1F1B, ScheduleGPipe, split_sequential

    torch.manual_seed(0)
    full = nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 16), nn.Tanh(), nn.Linear(16, 16), nn.Tanh(),
                         nn.Linear(16, 4))
    ref = nn.Sequential(*[m for m in full])
    import copy

    ref = copy.deepcopy(full)
    stage_mod = split_sequential(full, world, rank)
    stage = PipelineStage(stage_mod, rank, world, device=torch.device("cpu"))
    loss_fn = lambda out, tgt: ((out - tgt) ** 2).mean()
    sched = (Schedule1F1B if kind == "1f1b" else ScheduleGPipe)(stage, n_microbatches=4, loss_fn=loss_fn)
    g = torch.Generator().manual_seed(2)
    x, y = torch.randn(8, 8, generator=g), torch.randn(8, 4, generator=g)
    losses = []
    if rank == 0:
        sched.step(x)
    elif rank == world - 1:
        sched.step(target=y, losses=losses)
    else:
        sched.step()
    # reference: whole model, whole batch, mean of microbatch losses
    out = ref(x)
    l = sum(loss_fn(o, t) for o, t in zip(out.chunk(4), y.chunk(4))) / 4
    l.backward()
    mine = dict(stage_mod.named_parameters())
    refp = dict(split_sequential(ref, world, rank).named_parameters())
    for k, p in mine.items():
        assert torch.allclose(p.grad, refp[k].grad, atol=1e-6), (rank, k)
    if rank == world - 1:
        assert abs(float(sum(losses) / 4) - float(l)) < 1e-6


def test_pipeline_1f1b_grads_match():
    run_distributed(_pp_worker, 4, args=("1f1b",))


def test_pipeline_gpipe_grads_match():
    run_distributed(_pp_worker, 3, args=("gpipe",))


def _vp_xent_worker(rank, world):
    from b200dist import ops
    from b200dist.parallel import init_device_mesh
    from b200dist.parallel.tp import llama_tp_plan, vocab_parallel_cross_entropy

    ref = _toy_llama()
    model = _toy_llama()
    mesh = init_device_mesh("cpu", (world,), ("tp",))
    llama_tp_plan(model, mesh["tp"], sequence_parallel=True, gather_logits=False)
    tok = torch.randint(0, 64, (2, 9), generator=torch.Generator().manual_seed(5))
    x, y = tok[:, :-1], tok[:, 1:].clone()
    y[0, 2] = -100                                     # an ignored position
    out = model(x)
    assert out.shape == (2, 8, 64 // world)            # logits stay vocabulary-sharded: nothing is gathered
    loss = vocab_parallel_cross_entropy(out, y, mesh["tp"].communicator())
    loss_ref = ops.cross_entropy(ref(x), y)
    assert torch.allclose(loss, loss_ref, atol=1e-5), (float(loss), float(loss_ref))
    loss.backward(); loss_ref.backward()
    g = model.output.weight.grad
    gr = ref.output.weight.grad.chunk(world, 0)[rank]
    assert torch.allclose(g, gr, atol=1e-5, rtol=1e-3), (g - gr).abs().max()
    g = model.layers[0].attention.wq.weight.grad
    gr = ref.layers[0].attention.wq.weight.grad.chunk(world, 0)[rank]
    assert torch.allclose(g, gr, atol=1e-5, rtol=1e-3), (g - gr).abs().max()


def test_vocab_parallel_cross_entropy_matches_gathered_logits():
    run_distributed(_vp_xent_worker, 2)
