"""Provision the UNMODIFIED reference under ``baseline/_ref`` (git-ignored, ships to the GPU box).

    python baseline/install_ref.py            # idempotent; prints where the install lives
    from baseline.install_ref import ensure_reference, reference_path

Layout created (what every reference arm in this repo imports from):

    baseline/_ref/utils/...                   pip-installed package (the reference's pyproject packages ``utils*`` only)
    baseline/_ref/reference_src/fsdp_tp/      verbatim copies of the recipe trees that are NOT packaged:
    baseline/_ref/reference_src/scripts/      the Llama model (fsdp_tp/llama2_model.py), the recipes, the comm bench
    baseline/_ref/reference_src/tests/
    baseline/_ref/.installed                  marker: source path + how it was installed

The reference tree (``/root/reference`` or ``$B200DIST_REFERENCE``) is read-only, so pip builds from a
throw-away copy under the system temp directory.  If pip cannot run (no wheelhouse, no setuptools) the
package directory is copied instead - the reference is pure Python, the result is byte-identical to what
pip would lay down - and the marker records that.

``ensure_reference()`` never raises: it returns the install root or ``None`` (with the reason available from
``why_unavailable()``), so benches can print ``{"impl": "reference", "unavailable": ...}`` and GPU tests can skip.
Multi-process launches (torchrun) are safe: the install happens under an exclusive file lock and the other
ranks find the marker when they get the lock.
"""
from __future__ import annotations

import fcntl
import os
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_ROOT = HERE / "_ref"
MARKER = REF_ROOT / ".installed"
COPIED_TREES = ("fsdp_tp", "scripts", "tests")
_why = ""


def reference_source() -> Path | None:
    for cand in (os.environ.get("B200DIST_REFERENCE"), "/root/reference"):
        if cand and (Path(cand) / "fsdp_tp" / "llama2_model.py").exists():
            return Path(cand)
    return None


def is_installed() -> bool:
    return MARKER.exists() and (REF_ROOT / "reference_src" / "fsdp_tp" / "llama2_model.py").exists() \
        and (REF_ROOT / "utils" / "__init__.py").exists()


def why_unavailable() -> str:
    return _why


def _pip_install(src_copy: Path) -> tuple[bool, str]:
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--quiet",
           "--find-links", "/opt/wheelhouse", "--target", str(REF_ROOT), str(src_copy)]
    try:
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    except Exception as e:  # noqa: BLE001
        return False, f"{type(e).__name__}: {e}"
    return p.returncode == 0, (p.stderr or p.stdout)[-400:]


def _install(src: Path) -> str:
    REF_ROOT.mkdir(parents=True, exist_ok=True)
    with tempfile.TemporaryDirectory(prefix="refcopy_") as tmp:
        work = Path(tmp) / "reference"
        shutil.copytree(src, work, ignore=shutil.ignore_patterns(".git", "pixi.lock", "__pycache__"))
        ok, msg = _pip_install(work)
    how = "pip install --no-index --no-build-isolation --no-deps --target baseline/_ref <copy of reference>"
    if not ok or not (REF_ROOT / "utils" / "__init__.py").exists():
        # pure-Python package: copying it is what pip would have done
        shutil.copytree(src / "utils", REF_ROOT / "utils", dirs_exist_ok=True,
                        ignore=shutil.ignore_patterns("__pycache__"))
        how = f"copied utils/ (pip failed: {msg.strip().splitlines()[-1] if msg.strip() else 'unknown'})"
    for tree in COPIED_TREES:
        shutil.copytree(src / tree, REF_ROOT / "reference_src" / tree, dirs_exist_ok=True,
                        ignore=shutil.ignore_patterns("__pycache__"))
    MARKER.write_text(f"source: {src}\nhow: {how}\n")
    return how


def ensure_reference(verbose: bool = False) -> Path | None:
    """Install once (cross-process safe); returns baseline/_ref or None when no reference tree is reachable."""
    global _why
    if is_installed():
        return REF_ROOT
    src = reference_source()
    if src is None:
        _why = ("baseline/_ref is absent and no reference tree at /root/reference or $B200DIST_REFERENCE "
                "(run `python baseline/install_ref.py` where the reference is mounted)")
        return None
    HERE.mkdir(parents=True, exist_ok=True)
    try:
        with open(HERE / ".install.lock", "w") as lock:
            fcntl.flock(lock, fcntl.LOCK_EX)
            try:
                if not is_installed():
                    how = _install(src)
                    if verbose:
                        print(f"[install_ref] {REF_ROOT} <- {src}: {how}")
            finally:
                fcntl.flock(lock, fcntl.LOCK_UN)
    except Exception as e:  # noqa: BLE001
        _why = f"reference install failed: {type(e).__name__}: {str(e)[:160]}"
        return None
    if not is_installed():
        _why = "reference install did not produce baseline/_ref/reference_src/fsdp_tp/llama2_model.py"
        return None
    return REF_ROOT


def reference_path(*parts: str) -> Path:
    """Path inside the installed reference sources, e.g. reference_path('fsdp_tp')."""
    return REF_ROOT.joinpath("reference_src", *parts)


if __name__ == "__main__":
    root = ensure_reference(verbose=True)
    if root is None:
        print(f"[install_ref] unavailable: {why_unavailable()}")
        sys.exit(1)
    print(f"[install_ref] ok: {root}\n{MARKER.read_text()}", end="")
