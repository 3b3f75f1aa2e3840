"""Install the UNMODIFIED reference into ``baseline/_ref`` (git-ignored, ships with the working tree).

The documented command is
    python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref /root/reference
It fails on dependency resolution here (``torch>=2.0.0`` is not a wheel in the wheelhouse), so ``--no-deps`` is
added; the source tree is read-only, so pip builds from a copy under a temp dir.  If pip itself cannot build the
wheel (no setuptools backend, ...), the package directory is copied verbatim — the reference is pure Python, an
installed copy and a copied one are byte-identical (`pyproject.toml:74-75` only adds a console script).

``ensure_reference()`` is called by ``__graft_entry__.build()`` (runs where ``/root/reference`` is mounted) and
by ``reference_arm.run()`` before it gives up, so the reference arm does not depend on anyone remembering the
manual step.  Nothing in the reference is edited.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_DST = Path(os.environ["LLMCTL_REFERENCE_DST"]) if os.environ.get("LLMCTL_REFERENCE_DST") else HERE / "_ref"
MARK = "llmctl/runtime/engine.py"
SOURCES = [os.environ.get("LLMCTL_REFERENCE_SRC"), "/root/reference"]


def installed() -> bool:
    return (REF_DST / MARK).exists()


def _source() -> Path | None:
    for s in SOURCES:
        if s and (Path(s) / MARK).exists():
            return Path(s)
    return None


def _pip_install(src_copy: Path, dst: Path) -> tuple[bool, str]:
    base = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
            "/opt/wheelhouse", "--target", str(dst), "--upgrade", "--quiet"]
    log = []
    for extra in ([], ["--no-deps"]):
        r = subprocess.run(base + extra + [str(src_copy)], capture_output=True, text=True)
        log.append(f"pip {' '.join(extra) or '(with deps)'}: rc={r.returncode} {r.stderr.strip().splitlines()[-1:] or ''}")
        if r.returncode == 0 and (dst / MARK).exists():
            return True, "; ".join(log)
    return False, "; ".join(log)


def ensure_reference(verbose: bool = False) -> tuple[bool, str]:
    """Returns (available, how).  Safe to call from several ranks: one installs (lock dir), the rest wait."""
    if installed():
        return True, "already installed"
    src = _source()
    if src is None:
        return False, "baseline/_ref is empty and no reference source tree is mounted (/root/reference)"
    lock = REF_DST.parent / "._ref_lock"
    try:
        if lock.exists() and time.time() - lock.stat().st_mtime > 600:
            shutil.rmtree(lock, ignore_errors=True)      # left behind by an installer that died
    except OSError:
        pass
    try:
        lock.mkdir()
    except FileExistsError:
        for _ in range(600):                 # another rank is installing
            if installed() and not lock.exists():
                return True, "installed by another rank"
            time.sleep(0.5)
        return installed(), "waited for another rank's install"
    try:
        tmp = Path(tempfile.mkdtemp(prefix="llmctl_ref_src_"))
        copy = tmp / "reference"
        shutil.copytree(src, copy, ignore=shutil.ignore_patterns(".git", "__pycache__"))
        REF_DST.parent.mkdir(parents=True, exist_ok=True)
        stage = Path(tempfile.mkdtemp(prefix="_ref_stage_", dir=str(REF_DST.parent)))
        ok, how = _pip_install(copy, stage)
        if not ok:
            shutil.copytree(copy / "llmctl", stage / "llmctl", dirs_exist_ok=True)
            how += "; fell back to a verbatim copy of the pure-Python package"
        if REF_DST.exists():
            shutil.rmtree(REF_DST)
        stage.rename(REF_DST)
        shutil.rmtree(tmp, ignore_errors=True)
        if verbose:
            print(f"[reference] {how}")
        return installed(), how
    finally:
        shutil.rmtree(lock, ignore_errors=True)


if __name__ == "__main__":
    ok, how = ensure_reference(verbose=True)
    print(f"reference available: {ok} ({how})")
    sys.exit(0 if ok else 1)
