"""Fused optimizers with torch.optim semantics on hand-written kernels.

``FusedAdamW`` == torch.optim.AdamW (reference: 01_fsdp_tp_hybrid.py:162, multinode_ddp_unet.py),
``FusedSGD``  == torch.optim.SGD(momentum, weight_decay) (reference: scripts/main.py:309-311).
bf16 parameters get an fp32 master copy in the optimizer state; the kernel writes the master
weight and the bf16 compute copy in the same pass.  CPU parameters use the PyTorch math.
"""
from __future__ import annotations

import torch

from . import native as N


class FusedAdamW(torch.optim.Optimizer):
    """``capturable=True`` keeps the step-dependent scalars (lr, bias corrections) in a small DEVICE tensor per parameter
    group, refreshed by ``prepare_step()`` with a stream-ordered copy, and runs the ``adamw_dev`` kernel that reads them:
    ``step()`` can then be recorded into a CUDA graph once and replayed every iteration (``engine.Trainer(cuda_graph=True)``
    calls ``prepare_step()`` before each replay; learning-rate schedules keep working).  One step count per group."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, grad_scale=1.0,
                 foreach=None, fused=None, capturable: bool = False, overlap_backward: bool = False):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.grad_scale = grad_scale
        self.capturable = capturable
        self._external_prepare = False     # True while a Trainer drives prepare_step() itself (capture / replay)
        self._overlap_stream = None
        if overlap_backward:
            self.enable_backward_overlap()

    # ------------------------------------------------------------------ optimizer-in-backward
    def enable_backward_overlap(self) -> None:
        """Update each CUDA parameter as soon as autograd has finished its gradient, on a side stream, while the rest of
        backward keeps the tensor cores busy: the update is pure HBM traffic (28 bytes per parameter), the GEMMs it
        runs beside are compute bound, so on one GPU the optimizer pass (18 % of a Llama-7B step when it runs after
        backward) mostly disappears behind them - the single-GPU counterpart of what FSDP does with its fused
        reduce-scatter + AdamW per unit.  ``step()`` then only joins the side stream (and handles parameters that have
        no hook, e.g. CPU ones).  One backward == one optimizer step while this is on, so it does not combine with
        gradient accumulation or global-norm clipping; ``disable_backward_overlap()`` returns to the plain behaviour."""
        if self._overlap_stream is not None or self.capturable:
            return
        cuda_params = [(g, p) for g in self.param_groups for p in g["params"] if p.is_cuda and p.requires_grad]
        if not cuda_params:
            return
        self._overlap_stream = torch.cuda.Stream(cuda_params[0][1].device)
        self._overlap_hooks = [p.register_post_accumulate_grad_hook(self._make_overlap_hook(g)) for g, p in cuda_params]

    def disable_backward_overlap(self) -> None:
        if self._overlap_stream is None:
            return
        torch.cuda.current_stream().wait_stream(self._overlap_stream)
        for h in self._overlap_hooks:
            h.remove()
        self._overlap_stream, self._overlap_hooks = None, []

    def _make_overlap_hook(self, group):
        def hook(p):
            g = p.grad
            if g is None or self._overlap_stream is None:
                return
            cur = torch.cuda.current_stream(p.device)
            ev = torch.cuda.Event()
            ev.record(cur)
            with torch.cuda.stream(self._overlap_stream), torch.no_grad():
                self._overlap_stream.wait_event(ev)
                self._update(group, p, g)
            g.record_stream(self._overlap_stream)   # the caching allocator must not hand g out before the update read it
            p.grad = None
        return hook

    def _update(self, group, p, g) -> None:
        b1, b2 = group["betas"]
        st = self.state[p]
        if not st:
            st["step"] = 0
            st["exp_avg"] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
            st["exp_avg_sq"] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
            if p.dtype == torch.bfloat16:
                st["master"] = p.detach().float()
        st["step"] = group["_t"] if self.capturable else st["step"] + 1
        if (self.capturable and p.is_cuda and p.is_contiguous() and g.is_contiguous()
                and g.dtype in (torch.float32, torch.bfloat16)):
            master = st.get("master")
            N.adamw_step_dev(master if master is not None else p.data, g, st["exp_avg"], st["exp_avg_sq"],
                             group["_scalars"], beta1=b1, beta2=b2, eps=group["eps"], wd=group["weight_decay"],
                             grad_scale=self.grad_scale, p_lowp=p.data if master is not None else None)
        elif p.is_cuda and p.is_contiguous() and g.is_contiguous() and g.dtype in (torch.float32, torch.bfloat16):
            master = st.get("master")
            N.adamw_step(master if master is not None else p.data, g, st["exp_avg"], st["exp_avg_sq"],
                         lr=group["lr"], beta1=b1, beta2=b2, eps=group["eps"], wd=group["weight_decay"],
                         step=st["step"], grad_scale=self.grad_scale,
                         p_lowp=p.data if master is not None else None)
        else:
            gf = g.float() * self.grad_scale
            master = st.get("master", None)
            w = master if master is not None else p.data.float()
            st["exp_avg"].mul_(b1).add_(gf, alpha=1 - b1)
            st["exp_avg_sq"].mul_(b2).addcmul_(gf, gf, value=1 - b2)
            bc1, bc2 = 1 - b1 ** st["step"], 1 - b2 ** st["step"]
            w.mul_(1 - group["lr"] * group["weight_decay"])
            denom = (st["exp_avg_sq"].sqrt() / (bc2 ** 0.5)).add_(group["eps"])
            w.addcdiv_(st["exp_avg"], denom, value=-group["lr"] / bc1)
            p.data.copy_(w)

    def prepare_step(self) -> None:
        """Advance every group's step count and upload {lr, 1 - beta1^t, sqrt(1 - beta2^t)} to its device scalars."""
        for group in self.param_groups:
            dev = next((p.device for p in group["params"]), torch.device("cpu"))
            group["_t"] = group.get("_t", 0) + 1
            t, (b1, b2) = group["_t"], group["betas"]
            host = group.get("_scalars_host")
            if host is None:
                host = group["_scalars_host"] = torch.zeros(4, dtype=torch.float32, pin_memory=dev.type == "cuda")
                group["_scalars"] = torch.zeros(4, dtype=torch.float32, device=dev)
            host[0], host[1], host[2] = group["lr"], 1.0 - b1 ** t, (1.0 - b2 ** t) ** 0.5
            group["_scalars"].copy_(host, non_blocking=True)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        if self.capturable and not self._external_prepare:
            self.prepare_step()
        if self._overlap_stream is not None:   # updates issued from the gradient hooks: the next forward must see them
            torch.cuda.current_stream(self._overlap_stream.device).wait_stream(self._overlap_stream)
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None:
                    self._update(group, p, p.grad)
        return loss


class FusedSGD(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-2, momentum=0.0, weight_decay=0.0, grad_scale=1.0):
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay))
        self.grad_scale = grad_scale

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            mu, wd, lr = group["momentum"], group["weight_decay"], group["lr"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                first = not st
                if first:
                    st["step"] = 0
                    if mu != 0:
                        st["momentum_buffer"] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
                    if p.dtype == torch.bfloat16:
                        st["master"] = p.detach().float()
                st["step"] += 1
                g = p.grad
                buf = st.get("momentum_buffer")
                master = st.get("master")
                if p.is_cuda and p.is_contiguous() and g.is_contiguous() and g.dtype in (torch.float32, torch.bfloat16):
                    N.sgd_step(master if master is not None else p.data, g, buf, lr=lr, momentum=mu, wd=wd,
                               first_step=first, grad_scale=self.grad_scale,
                               p_lowp=p.data if master is not None else None)
                else:
                    w = master if master is not None else p.data.float()
                    d = g.float() * self.grad_scale + wd * w
                    if buf is not None:
                        if first:
                            buf.copy_(d)
                        else:
                            buf.mul_(mu).add_(d)
                        d = buf
                    w.add_(d, alpha=-lr)
                    p.data.copy_(w)
        return loss
