"""One config dataclass for every entrypoint: the reference's 13 ``TrainingConfig`` fields
(/root/reference/utils/config.py:25-122, same names, same ``--lr`` alias, YAML loader that
ignores unknown keys) plus the knobs the reference hard-codes (tp/dp/pp sizes, model preset,
sequence length, dtype, microbatches, schedule: SURVEY.md §5.6)."""
from __future__ import annotations

import argparse
import dataclasses
from dataclasses import dataclass, fields
from typing import Any, Dict, Optional, Sequence


@dataclass
class TrainingConfig:
    # --- reference fields (utils/config.py:30-50) ---
    num_epochs: int = 10
    batch_size: int = 32
    learning_rate: float = 1e-3
    weight_decay: float = 0.0
    momentum: float = 0.9
    seed: int = 42
    backend: str = "nccl"
    use_amp: bool = False
    amp_dtype: str = "bfloat16"
    save_every: int = 0
    checkpoint_dir: str = "./checkpoints"
    profile: bool = False
    profile_dir: str = "./profiler_output"
    # --- additions (hard-coded in the reference recipes) ---
    model: str = "llama2-toy"
    seq_len: int = 256
    dtype: str = "bfloat16"
    tp: int = 1
    dp: int = 0            # 0 = world_size / (tp * pp)
    pp: int = 1
    sp: bool = True        # Megatron sequence parallelism inside the TP group
    microbatches: int = 4
    schedule: str = "1f1b"
    steps: int = 10
    optimizer: str = "adamw"
    comm_ctas: int = 16    # persistent CTAs reserved for fused collectives
    oneshot_threshold: int = 262144  # bytes; below this all-reduce is one-shot

    def to_dict(self) -> Dict[str, Any]:
        return dataclasses.asdict(self)

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "TrainingConfig":
        names = {f.name for f in fields(cls)}
        return cls(**{k: v for k, v in d.items() if k in names})

    @classmethod
    def from_yaml(cls, path: str) -> "TrainingConfig":
        import yaml

        with open(path) as f:
            raw = yaml.safe_load(f) or {}
        return cls.from_dict(raw)

    @classmethod
    def add_arguments(cls, parser: argparse.ArgumentParser) -> argparse.ArgumentParser:
        d = cls()
        p = parser
        p.add_argument("--num_epochs", "--num-epochs", "--epochs", type=int, default=d.num_epochs)
        p.add_argument("--batch_size", "--batch-size", type=int, default=d.batch_size)
        p.add_argument("--learning_rate", "--learning-rate", "--lr", type=float, default=d.learning_rate)
        p.add_argument("--weight_decay", "--weight-decay", type=float, default=d.weight_decay)
        p.add_argument("--momentum", type=float, default=d.momentum)
        p.add_argument("--seed", "--random_seed", type=int, default=d.seed)
        p.add_argument("--backend", type=str, default=d.backend, choices=["nccl", "gloo", "mpi"])
        p.add_argument("--use_amp", "--use-amp", action="store_true", default=d.use_amp)
        p.add_argument("--amp_dtype", "--amp-dtype", type=str, default=d.amp_dtype,
                       choices=["bfloat16", "float16"])
        p.add_argument("--save_every", "--save-every", type=int, default=d.save_every)
        p.add_argument("--checkpoint_dir", "--checkpoint-dir", type=str, default=d.checkpoint_dir)
        p.add_argument("--profile", action="store_true", default=d.profile)
        p.add_argument("--profile_dir", "--profile-dir", type=str, default=d.profile_dir)
        p.add_argument("--model", type=str, default=d.model)
        p.add_argument("--seq_len", "--seq-len", type=int, default=d.seq_len)
        p.add_argument("--dtype", type=str, default=d.dtype, choices=["bfloat16", "float32"])
        p.add_argument("--tp", "--tp_size", type=int, default=d.tp)
        p.add_argument("--dp", "--fsdp_size", type=int, default=d.dp)
        p.add_argument("--pp", type=int, default=d.pp)
        p.add_argument("--no-sp", dest="sp", action="store_false", default=d.sp)
        p.add_argument("--microbatches", "--num-microbatches", type=int, default=d.microbatches)
        p.add_argument("--schedule", type=str, default=d.schedule, choices=["gpipe", "1f1b"])
        p.add_argument("--steps", "--num-steps", type=int, default=d.steps)
        p.add_argument("--optimizer", type=str, default=d.optimizer, choices=["adamw", "sgd"])
        p.add_argument("--comm-ctas", dest="comm_ctas", type=int, default=d.comm_ctas)
        p.add_argument("--oneshot-threshold", dest="oneshot_threshold", type=int, default=d.oneshot_threshold)
        p.add_argument("--config", type=str, default=None, help="YAML file; CLI flags override it")
        return p

    @classmethod
    def from_args(cls, args: Optional[Sequence[str]] = None) -> "TrainingConfig":
        """Parse known flags only (unknown flags are left for the caller, like the reference)."""
        parser = argparse.ArgumentParser(add_help=False)
        cls.add_arguments(parser)
        ns, _ = parser.parse_known_args(args)
        base = cls.from_yaml(ns.config) if ns.config else cls()
        defaults = cls()
        out = base.to_dict()
        for f in fields(cls):
            v = getattr(ns, f.name, None)
            if v is not None and v != getattr(defaults, f.name):
                out[f.name] = v
        return cls(**out)

    def torch_dtype(self):
        import torch

        return {"bfloat16": torch.bfloat16, "float32": torch.float32, "float16": torch.float16}[self.dtype]
