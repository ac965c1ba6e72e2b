"""Model-level CPU tests: initialisation statistics, pipeline-stage slicing of the Llama model."""
import torch

from b200dist.models import ModelArgs, Transformer, TransformerStage

CFG = dict(dim=128, n_layers=4, n_heads=4, vocab_size=512, max_seq_len=64)


def test_bf16_init_has_the_requested_std():
    # trunc_normal_ evaluated in bf16 gives std ~0.13 for a requested 0.02; the model must draw in fp32 and round once
    prev = torch.get_default_dtype()
    try:
        torch.set_default_dtype(torch.bfloat16)
        torch.manual_seed(0)
        m = Transformer(ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=1024, max_seq_len=64))
    finally:
        torch.set_default_dtype(prev)
    assert m.layers[0].attention.wq.weight.dtype == torch.bfloat16
    assert abs(float(m.layers[0].attention.wq.weight.float().std()) - 0.02) < 2e-3
    assert abs(float(m.layers[1].feed_forward.w2.weight.float().std()) - 0.02 / 4 ** 0.5) < 2e-3
    assert abs(float(m.output.weight.float().std()) - 256 ** -0.5) < 6e-3
    assert m.freqs_cis.dtype == torch.complex64


def test_stage_slices_compose_to_the_full_model():
    torch.manual_seed(0)
    full = Transformer(ModelArgs(**CFG))
    sd = full.state_dict()
    stages = [TransformerStage(ModelArgs(**CFG), i, 2, init=False) for i in range(2)]
    keys = set()
    for st in stages:
        own = st.state_dict()
        keys |= set(own)
        st.load_state_dict({k: sd[k] for k in own})
    assert keys == set(sd) - {"freqs_cis"}          # union of the stage state dicts == full model (RoPE table is derived)
    assert stages[0].layer_ids == [0, 1] and stages[1].layer_ids == [2, 3]
    tok = torch.randint(0, 512, (2, 33))
    h = stages[0](tok)
    assert h.shape == (2, 33, 128)
    out = stages[1](h)
    torch.testing.assert_close(out, full(tok), rtol=1e-5, atol=1e-5)
    # uneven split: the last stage takes the remainder
    odd = [TransformerStage(ModelArgs(**{**CFG, "n_layers": 5}), i, 2, init=False).layer_ids for i in range(2)]
    assert odd == [[0, 1, 2], [3, 4]]


def test_pack_block_weights_makes_qkv_and_w13_adjacent_and_keeps_values():
    """models/fused_block.py: packing re-homes q|k|v and w1|w3 into contiguous buffers (values unchanged) so the fused
    block node can address each group as one row-stacked matrix."""
    import torch

    from b200dist.models import fused_block
    from b200dist.models.llama2 import ModelArgs, Transformer

    torch.manual_seed(0)
    m = Transformer(ModelArgs(dim=32, n_layers=1, n_heads=4, vocab_size=64, max_seq_len=16, multiple_of=16))
    blk = m.layers[0]
    before = {n: p.detach().clone() for n, p in blk.named_parameters()}
    assert fused_block.adjacent_rows((blk.attention.wq.weight, blk.attention.wk.weight, blk.attention.wv.weight)) is None
    fused_block.pack_block_weights(blk)
    qkv = fused_block.adjacent_rows((blk.attention.wq.weight, blk.attention.wk.weight, blk.attention.wv.weight))
    w13 = fused_block.adjacent_rows((blk.feed_forward.w1.weight, blk.feed_forward.w3.weight))
    assert qkv is not None and tuple(qkv.shape) == (96, 32) and w13 is not None
    assert torch.equal(qkv[32:64], blk.attention.wk.weight) and torch.equal(w13[w13.shape[0] // 2:], blk.feed_forward.w3.weight)
    for n, p in blk.named_parameters():
        assert torch.equal(p, before[n]), n
    # an in-place update of one parameter is visible through the stacked view (same storage)
    with torch.no_grad():
        blk.attention.wv.weight.add_(1.0)
    assert torch.equal(qkv[64:], blk.attention.wv.weight)
    # CPU tensors keep the module-by-module path: the fused node never runs here
    x = torch.randn(1, 8, 32)
    assert not fused_block.block_supported(blk, x)
    y = m(torch.randint(0, 64, (1, 8)))
    assert y.shape == (1, 8, 64)


def test_fsdp_unit_layout_pulls_adjacent_groups_together():
    from b200dist.models.llama2 import ModelArgs, Transformer
    from b200dist.parallel.fsdp import _ordered_named_parameters

    m = Transformer(ModelArgs(dim=32, n_layers=1, n_heads=4, vocab_size=64, max_seq_len=16, multiple_of=16))
    names = [n for n, _ in _ordered_named_parameters(m.layers[0])]
    i = names.index("attention.wq.weight")
    assert names[i:i + 3] == ["attention.wq.weight", "attention.wk.weight", "attention.wv.weight"]
    j = names.index("feed_forward.w1.weight")
    assert names[j + 1] == "feed_forward.w3.weight"
    assert sorted(names) == sorted(n for n, _ in m.layers[0].named_parameters())


def test_attention_probability_dropout_and_host_philox_mask():
    """nn.MultiheadAttention(dropout=p) semantics of the pipeline transformer block (reference
    scripts/04_pipeline_parallel_pp/03_pipeline_training.py:57-58): dropout acts on the attention probabilities in
    training mode only; the host Philox restatement (oracle of the CUDA kernel) is deterministic in (seed, offset)."""
    from b200dist import ops
    from b200dist.models.pp_transformer import Block
    from b200dist.ops.functional import philox_keep_mask

    m1 = philox_keep_mask(40000, 0.25, seed=77, offset=3)
    assert torch.equal(m1, philox_keep_mask(40000, 0.25, seed=77, offset=3))
    assert not torch.equal(m1, philox_keep_mask(40000, 0.25, seed=77, offset=4))
    assert not torch.equal(m1, philox_keep_mask(40000, 0.25, seed=78, offset=3))
    assert abs(m1.float().mean().item() - 0.75) < 0.02
    assert torch.equal(m1[:1000], philox_keep_mask(1000, 0.25, seed=77, offset=3))      # counter-based: a prefix is a prefix
    assert philox_keep_mask(1000, 0.0, 1).all()

    torch.manual_seed(0)
    q, k, v = (torch.randn(2, 12, 4, 16) for _ in range(3))
    plain = ops.attention(q, k, v, causal=False)
    assert torch.allclose(ops.attention_dropout(q, k, v, 0.3, training=False), plain)
    assert torch.allclose(ops.attention_dropout(q, k, v, 0.0, training=True), plain)
    dropped = ops.attention_dropout(q, k, v, 0.3, training=True)
    assert dropped.shape == plain.shape and not torch.allclose(dropped, plain)

    blk = Block(32, 4, dropout=0.2)
    x = torch.randn(2, 12, 32, requires_grad=True)
    blk.train()
    y1, y2 = blk(x), blk(x)
    assert not torch.equal(y1, y2)
    y1.sum().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all()
    blk.eval()
    assert torch.equal(blk(x), blk(x))
