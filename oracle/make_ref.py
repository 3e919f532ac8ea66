"""Recipe: install the UNMODIFIED reference package where it can travel to the GPU box.

TEST INFRASTRUCTURE ONLY.  Reference sources are never copied into the repository's history: this
script pip-installs `/root/reference` (pure Python, `pyproject.toml:1-27`, no native build) with the
offline recipe of the task contract

    python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
        --target <dst> <copy of /root/reference under /tmp>

(`--no-deps`: torch / triton / transformers / flash-attn / xxhash are already in the image; a copy,
because setuptools writes `build/` and `*.egg-info` into the source tree and `/root/reference` is
read-only) into two git-ignored, NOT gpurun-ignored directories:

  oracle/_ref/     the checker: `oracle/reference_live.py` imports `nanovllm.models.glm4_moe.moe`
                   from here to pin the restatement on the GPU box and to time the reference's own
                   CPU torch path (`bench.py --impl reference`, `cpu_baseline.kind = "reference"`)
  baseline/_ref/   the host application for BASELINE configs[3]: the reference's engine / decoder
                   layer / attention calling OUR sparse block (`glm4_moe_b200.engine_shim`)

`__graft_entry__.build()` runs it whenever `/root/reference` is present (this container); the GPU box
only uses the installed files.  Usage:  python oracle/make_ref.py [--force]
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = "/root/reference"
TARGETS = (os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "baseline", "_ref"))
MARKER = os.path.join("nanovllm", "models", "glm4_moe", "moe.py")


def installed(target: str) -> bool:
    return os.path.exists(os.path.join(target, MARKER))


def ensure_ref(force: bool = False, verbose: bool = False) -> list[str]:
    """Install the reference into every target that lacks it.  Returns the targets that hold it."""
    todo = [t for t in TARGETS if force or not installed(t)]
    if todo and os.path.isdir(os.path.join(REFERENCE, "nanovllm")):
        with tempfile.TemporaryDirectory(prefix="glm4moe_ref_") as tmp:
            src = os.path.join(tmp, "reference")
            shutil.copytree(REFERENCE, src, ignore=shutil.ignore_patterns(".git", "__pycache__"))
            for t in todo:
                if os.path.isdir(t):
                    shutil.rmtree(t)
                os.makedirs(os.path.dirname(t), exist_ok=True)
                cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps",
                       "--find-links", "/opt/wheelhouse", "--target", t, src]
                r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
                if r.returncode != 0:
                    raise RuntimeError(f"pip install of the reference into {t} failed:\n{r.stdout}")
                if verbose:
                    print(r.stdout.strip().splitlines()[-1])
    return [t for t in TARGETS if installed(t)]


if __name__ == "__main__":
    have = ensure_ref(force="--force" in sys.argv, verbose=True)
    print("reference installed in:", ", ".join(os.path.relpath(t, ROOT) for t in have) or "(nowhere: no /root/reference)")
