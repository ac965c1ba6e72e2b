"""In-tree build of the native library ``b200dist/_native/libb200dist.so``.

Every ``csrc/*.cu`` / ``csrc/*.cpp`` file is compiled for sm_100a only
(``-gencode arch=compute_100a,code=sm_100a -lineinfo``) and linked into ONE shared
object with a plain C ABI (declared in ``include/b200dist.h``).  The library is
loaded with ctypes (``b200dist.ops._lib``), so it does not link against libtorch
and nvcc cross-compiles it on a box without a GPU.

Run as ``python -m b200dist._build [--force] [--jobs N]``.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OUT_DIR = ROOT / "b200dist" / "_native"
OBJ_DIR = ROOT / "build" / "obj"
LIB = OUT_DIR / "libb200dist.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr", "--use_fast_math",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-O3",
    "-Xptxas", "-v",
    "-I", str(CSRC), "-I", str(ROOT / "include"),
]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC or install the CUDA toolkit")


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((ROOT / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()


def _compile_one(nvcc: str, src: Path, force: bool) -> tuple[Path, str]:
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha")
    dig = _digest(src)
    if not force and obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj, ""
    cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    (OBJ_DIR / (src.stem + ".ptxas.log")).write_text(r.stderr)
    return obj, r.stderr


def build(force: bool = False, jobs: int | None = None, verbose: bool = False) -> Path:
    nvcc = nvcc_path()
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")))
    if not srcs:
        raise RuntimeError(f"no sources under {CSRC}")
    jobs = jobs or min(len(srcs), os.cpu_count() or 4)
    objs: list[Path] = []
    rebuilt = False
    with cf.ThreadPoolExecutor(max_workers=jobs) as ex:
        futs = {ex.submit(_compile_one, nvcc, s, force): s for s in srcs}
        for f in cf.as_completed(futs):
            obj, log = f.result()
            objs.append(obj)
            if log:
                rebuilt = True
                if verbose:
                    print(f"[b200dist build] {futs[f].name}\n{log}")
    if rebuilt or not LIB.exists():
        cmd = [nvcc, "-shared", "-o", str(LIB), *[str(o) for o in sorted(objs)],
               "-gencode", "arch=compute_100a,code=sm_100a", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--jobs", type=int, default=None)
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args()
    lib = build(a.force, a.jobs, a.verbose)
    print(f"built {lib} ({lib.stat().st_size / 1e6:.1f} MB)")


if __name__ == "__main__":
    sys.exit(main())
