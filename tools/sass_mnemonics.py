"""Blackwell-native SASS mnemonics per kernel (no GPU needed): cuobjdump -sass of the built library, demangled kernel names,
counts of tcgen05 / TMEM / TMA / multimem instructions.  python tools/sass_mnemonics.py > profiles/sass_mnemonics_r1.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

SO = Path(__file__).resolve().parent.parent / "b200dist" / "_native" / "libb200dist.so"
PAT = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UBLKCP[.\w]*|UTCBAR[.\w]*|UTCATOMSWS[.\w]*|"
                 r"LDGMC[.\w]*|REDG[.\w]*|RED\.E[.\w]*|HMMA[.\w]*|FFMA2|FADD2|FMUL2|FMNMX3|MUFU\.EX2)\b")
sass = subprocess.run(["cuobjdump", "-sass", str(SO)], capture_output=True, text=True, check=True).stdout
kern, counts = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        continue
    if kern:
        for tok in PAT.findall(line):
            counts[kern][tok] += 1
names = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
print("# cuobjdump -sass b200dist/_native/libb200dist.so : Blackwell-native mnemonics per kernel (UTC*MMA = tcgen05.mma, LDTM/STTM = "
      "tcgen05.ld/st,\n# UTMALDG = TMA tensor load (.4D also feeds the implicit-GEMM convolution), UBLKCP = TMA bulk copy, UTCBAR = "
      "tcgen05.commit, LDGMC = multimem.ld_reduce\n# (NVSwitch in-fabric reduction), FFMA2/FADD2/FMUL2 = packed fp32x2 math; no HMMA "
      "anywhere = no legacy mma.sync path)")
rows = []
for mangled, name in zip(counts, names):
    c = counts[mangled]
    if not c:
        continue
    name = re.sub(r"^void b200::", "", name)
    name = re.sub(r"\(.*$", "", name)
    rows.append((name, "  ".join(f"{k}x{v}" for k, v in sorted(c.items()))))
for name, txt in sorted(rows):
    print(f"{name:90s} {txt}")
print(f"# kernels with HMMA (legacy tensor-core path): {sum(1 for c in counts.values() if any(k.startswith('HMMA') for k in c))}",
      file=sys.stdout)
