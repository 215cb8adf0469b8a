"""Puts the UNMODIFIED reference under ``baseline/_ref`` (git-ignored, shipped to the GPU box with the snapshot).

Used by ``__graft_entry__.build()`` and by ``bench.py --impl reference``.  Order of attempts:

1. ``baseline/_ref/scaling`` already there -> nothing to do;
2. pip install (offline, ``--no-deps``: the reference pins torch==2.1.1 / flash-attn==2.4.2, which no offline wheel
   satisfies) from a scratch copy of the reference checkout, because its setup.py writes into the source tree;
3. plain copy of ``src/scaling`` (the reference is pure Python, nothing is compiled).

The source checkout is looked up in ``$SCALING_REFERENCE_SRC``, ``/root/reference`` and ``baseline/reference_src``.
Returns ``(ok, message)``; never raises.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_ROOT = HERE / "_ref"


def _sources() -> list[Path]:
    cands = [os.environ.get("SCALING_REFERENCE_SRC"), "/root/reference", str(HERE / "reference_src")]
    return [Path(c) for c in cands if c and (Path(c) / "src" / "scaling" / "__init__.py").is_file()]


def ensure_reference(verbose: bool = False) -> tuple[bool, str]:
    if (REF_ROOT / "scaling" / "__init__.py").is_file():
        return True, "present"
    srcs = _sources()
    if not srcs:
        return False, "baseline/_ref is missing and no reference checkout is reachable from this machine"
    src = srcs[0]
    errors = []
    try:
        with tempfile.TemporaryDirectory(prefix="refcopy_") as tmp:
            work = Path(tmp) / "reference"
            shutil.copytree(src, work, ignore=shutil.ignore_patterns(".git", "__pycache__"))
            cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps",
                   "--find-links", "/opt/wheelhouse", "--target", str(REF_ROOT), str(work)]
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
            if verbose:
                print(r.stdout[-2000:], r.stderr[-2000:])
            if r.returncode == 0 and (REF_ROOT / "scaling" / "__init__.py").is_file():
                return True, f"pip --no-deps --target baseline/_ref from {src}"
            errors.append(f"pip rc={r.returncode}: {r.stderr.strip().splitlines()[-1:]}")
    except Exception as e:  # noqa: BLE001
        errors.append(f"pip: {type(e).__name__}: {e}")
    try:
        REF_ROOT.mkdir(parents=True, exist_ok=True)
        shutil.copytree(src / "src" / "scaling", REF_ROOT / "scaling", dirs_exist_ok=True,
                        ignore=shutil.ignore_patterns("__pycache__"))
        return True, f"plain copy of {src}/src/scaling ({'; '.join(errors)})"
    except Exception as e:  # noqa: BLE001
        errors.append(f"copy: {type(e).__name__}: {e}")
    return False, "; ".join(errors)[:280]


if __name__ == "__main__":
    ok, msg = ensure_reference(verbose="-v" in sys.argv)
    print(("ok: " if ok else "FAILED: ") + msg)
    sys.exit(0 if ok else 1)
