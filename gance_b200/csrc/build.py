"""
Builds gance_b200/csrc/libgance_b200.so in-tree with nvcc for sm_100a (cross-compiles without a
GPU). The .so is git-ignored but travels to the GPU box with the repo snapshot.
"""

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path
from typing import List

CSRC = Path(__file__).resolve().parent
SOURCES = ["runtime.cu", "synthesis_kernels.cu", "modconv_tc.cu", "modconv_tf32.cu", "upconv_phase.cu", "upconv_fused.cu", "upconv_fused_pair.cu", "audio_kernels.cu", "resize_kernels.cu"]
# Measurement builds, never the product: GANCE_PIPE_STATS=1 -> libgance_b200_stats.so (cycle counters per warp role,
# tools/pipe_stats.py); GANCE_VARIANT=name:MACRO[,MACRO...] -> libgance_b200_<name>.so with -DMACRO... (ablations of the
# epilogues, tools/ablate.sh).
PIPE_STATS = os.environ.get("GANCE_PIPE_STATS") == "1"
VARIANT, _, VARIANT_MACROS = os.environ.get("GANCE_VARIANT", "").partition(":")
SUFFIX = "_stats" if PIPE_STATS else (f"_{VARIANT}" if VARIANT else "")
LIBRARY = CSRC / f"libgance_b200{SUFFIX}.so"
STAMP = CSRC / f"libgance_b200{SUFFIX}.stamp"

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-O3",
    "-std=c++17",
    "-lineinfo",
    "-Xcompiler",
    "-fPIC",
] + (["-DGANCE_PIPE_STATS"] if PIPE_STATS else []) + [f"-D{macro}" for macro in VARIANT_MACROS.split(",") if macro]


def _nvcc() -> str:
    candidates = [os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"]
    for candidate in candidates:
        if candidate and Path(candidate).exists():
            return candidate
    raise RuntimeError("nvcc not found; set NVCC or put /usr/local/cuda/bin on PATH")


def _existing_sources() -> List[Path]:
    return [CSRC / name for name in SOURCES if (CSRC / name).exists()]


def _fingerprint() -> str:
    digest = hashlib.sha256()
    files = sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [CSRC.parents[1] / "include" / "gance_b200.h"])
    for path in files:
        digest.update(path.name.encode())
        digest.update(path.read_bytes())
    digest.update(" ".join(NVCC_FLAGS).encode())
    return digest.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu into one shared library. Skips the work when sources are unchanged."""
    fingerprint = _fingerprint()
    if not force and LIBRARY.exists() and STAMP.exists() and STAMP.read_text().strip() == fingerprint:
        return LIBRARY
    nvcc = _nvcc()
    objects = []
    build_dir = CSRC / f"build{SUFFIX}"
    build_dir.mkdir(exist_ok=True)
    procs = []
    for source in _existing_sources():
        obj = build_dir / (source.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-Xptxas", "-v", "-c", str(source), "-o", str(obj)]
        procs.append((source, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for source, obj, proc in procs:
        output, _ = proc.communicate()
        (build_dir / (source.stem + ".ptxas.log")).write_text(output)
        if proc.returncode != 0:
            sys.stderr.write(output)
            raise RuntimeError(f"nvcc failed on {source.name}")
        if verbose:
            sys.stderr.write(output)
        objects.append(str(obj))
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIBRARY), *objects]
    result = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, check=False)
    if result.returncode != 0:
        sys.stderr.write(result.stdout)
        raise RuntimeError("nvcc link failed")
    STAMP.write_text(fingerprint)
    return LIBRARY


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
