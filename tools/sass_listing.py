"""Full SASS listings of the fused / hot kernels (no GPU needed): ``cuobjdump -sass`` of the built library, one file per
kernel under profiles/sass/, gzip-compressed, plus an index with instruction counts and the Blackwell-specific mnemonics.

    python tools/sass_listing.py                 # the default kernel set (fused compute+collective kernels first)
    python tools/sass_listing.py 'flash_attn'    # any kernels whose demangled name matches the regex

What to look for (B200_PROFILING.md): UTCHMMA(.2CTA) = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG = TMA tensor
load, UTMASTG = TMA tensor store, UBLKCP = TMA bulk copy (the peer pulls of the all-gather GEMM), UTCBAR = tcgen05.commit,
LDGMC = multimem.ld_reduce (NVSwitch in-fabric reduction); multimem.st is an ordinary STG.E.128.STRONG.SYS on the multicast
address; ST / LD ... .STRONG.SYS = peer-memory accesses with system scope.
"""
from __future__ import annotations

import gzip
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
SO = ROOT / "b200dist" / "_native" / "libb200dist.so"
OUT = ROOT / "profiles" / "sass"

DEFAULT = (r"gemm_tcgen05_kernel<2, 256, false, false, __nv_bfloat16, true, [01]>|"  # all-gather -> GEMM / GEMM -> reduce-scatter (1 = TMA-store epilogue)
           r"gemm_tcgen05_kernel<2, 256, false, true, __nv_bfloat16, true, [01]>|"
           r"gemm_tcgen05_kernel<2, 256, false, false, __nv_bfloat16, false, 0>|"   # plain forward GEMM (reference point)
           r"gemm_tcgen05_kernel<1, (64|128), false, false, __nv_bfloat16, true, 0>|"  # implicit-GEMM convolution fprop / dgrad / stem
           r"gemm_tcgen05_kernel<1, 128, true, true, float, true, 0>|"              # implicit-GEMM weight gradient
           r"comm_rs_adamw_kernel<__nv_bfloat16|nvls_rs_adamw_kernel<__nv_bfloat16|nvls_allreduce_kernel<__nv_bfloat16|"
           r"nvls_reducescatter_kernel<__nv_bfloat16, float|nvls_reducescatter_kernel<float, float|nvls_allgather_kernel|"
           r"nvls_broadcast_kernel|comm_slab_reduce_kernel|comm_allreduce_push_kernel<__nv_bfloat16, __nv_bfloat16, [01]>|"
           r"comm_gather_push_kernel|comm_neighbor_kernel|nvls_sync_kernel|"
           r"rmsnorm_bwd_kernel<__nv_bfloat16, 1>|rmsnorm_fwd_kernel<__nv_bfloat16>|softmax_dropout_fwd_kernel<__nv_bfloat16>|"
           r"flash_attn_fwd_kernel<128>|flash_attn_bwd_dkv_kernel<128>|flash_attn_bwd_dq_kernel<128>")
KEY = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UBLKCP[.\w]*|UTCBAR[.\w]*|LDGMC[.\w]*|"
                 r"STG\.E\.128\.STRONG\.SYS|LDG\.E\.128\.STRONG\.SYS|LD\.E\.128\.STRONG\.SYS|ST\.E\.128\.STRONG\.SYS|REDG[.\w]*|RED\.E[.\w]*|HMMA[.\w]*|MUFU\.EX2|FFMA2|SYNCS[.\w]*)")


def main():
    pat = re.compile(sys.argv[1] if len(sys.argv) > 1 else DEFAULT)
    sass = subprocess.run(["cuobjdump", "-sass", str(SO)], capture_output=True, text=True, check=True).stdout
    blocks, cur, name = [], [], None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name is not None:
                blocks.append((name, cur))
            name, cur = m.group(1), []
        if name is not None:
            cur.append(line)
    if name is not None:
        blocks.append((name, cur))
    dem = subprocess.run(["c++filt"], input="\n".join(b[0] for b in blocks), capture_output=True, text=True).stdout.splitlines()
    OUT.mkdir(parents=True, exist_ok=True)
    index = ["# profiles/sass/: cuobjdump -sass listings (gzip) of the fused / hot kernels, built from this tree",
             "# file | kernel | SASS instructions | Blackwell / multi-GPU mnemonics"]
    for (mangled, lines), full in zip(blocks, dem):
        short = re.sub(r"\(.*$", "", re.sub(r"^void b200::", "", full))
        if not pat.search(short):
            continue
        fname = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_")[:100] + ".sass.gz"
        with gzip.open(OUT / fname, "wt") as f:
            f.write(f"// {full}\n" + "\n".join(lines) + "\n")
        n_ins = sum(1 for l in lines if re.search(r"/\*[0-9a-f]{4}\*/", l))
        cnt = {}
        for l in lines:
            for tok in KEY.findall(l):
                cnt[tok] = cnt.get(tok, 0) + 1
        index.append(f"{fname} | {short} | {n_ins} | " + " ".join(f"{k}x{v}" for k, v in sorted(cnt.items())))
    (OUT / "INDEX.txt").write_text("\n".join(index) + "\n")
    print("\n".join(index))


if __name__ == "__main__":
    main()
