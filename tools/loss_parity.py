"""Loss-curve parity of the Llama model + optimizers against the reference model on ONE GPU (repeated batches so that
memorisation dynamics, which depend on every gradient being right, are visible).  python tools/loss_parity.py"""
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from baseline.install_ref import ensure_reference, reference_path, why_unavailable  # noqa: E402

if ensure_reference() is None:
    raise SystemExit(f"loss_parity needs the reference model: {why_unavailable()}")
sys.path.insert(0, str(reference_path("fsdp_tp")))
from llama2_model import ModelArgs as RA, Transformer as RT  # noqa: E402

from b200dist import ops  # noqa: E402
from b200dist.models import ModelArgs, Transformer  # noqa: E402
from b200dist.ops import FusedAdamW  # noqa: E402

dev = torch.device("cuda", 0)
dim = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
cfg = dict(dim=dim, n_layers=4, n_heads=dim // 128, vocab_size=32000, max_seq_len=2048)
S = 2048
g = torch.Generator().manual_seed(7)
toks = [torch.randint(0, 32000, (2, S + 1), generator=g).to(dev) for _ in range(2)]
torch.manual_seed(0)
with torch.device(dev):
    ref = RT(RA(**cfg))
sd = {k: v.clone() for k, v in ref.state_dict().items()}


def run(name, model, opt, loss_fn):
    ls = []
    for i in range(8):
        t = toks[i % 2]
        opt.zero_grad()
        loss = loss_fn(model(t[:, :-1]).view(-1, 32000), t[:, 1:].reshape(-1))
        loss.backward(); opt.step(); ls.append(round(float(loss), 3))
    print(f"{name:34s}", ls, flush=True)


kw = dict(lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1)
run("reference fp32 + torch AdamW", ref, torch.optim.AdamW(ref.parameters(), **kw), F.cross_entropy)
ref.load_state_dict(sd); ref = ref.to(torch.bfloat16)
run("reference bf16 + torch AdamW", ref, torch.optim.AdamW(ref.parameters(), **kw), F.cross_entropy)
del ref
for oname in ("torch AdamW", "FusedAdamW"):
    with torch.device(dev):
        m = Transformer(ModelArgs(**cfg))
    m.load_state_dict(sd); m = m.to(torch.bfloat16)
    opt = torch.optim.AdamW(m.parameters(), **kw) if oname == "torch AdamW" else FusedAdamW(m.parameters(), **kw)
    run(f"b200dist bf16 + {oname}", m, opt, ops.cross_entropy)
    del m, opt
