"""Build libb200llm.so (CUDA kernels + C ABI) in-tree for sm_100a.

    python build.py            # incremental
    python build.py --force    # rebuild everything

nvcc cross-compiles without a GPU. The shared library lands next to this file so it travels
with the source tree; nothing is installed into site-packages.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
BUILD = HERE / "build"
LIB = HERE / "libb200llm.so"
SOURCES = ["gemm_tcgen05.cu", "ops.cu", "attention.cu", "attention_tc.cu", "decode_fused.cu", "engine.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _stale(obj: Path, deps: list[Path]) -> bool:
    if not obj.exists():
        return True
    t = obj.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def _compile(src: Path, obj: Path) -> str:
    cmd = [NVCC, *FLAGS, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = (r.stdout or "") + (r.stderr or "")
    (BUILD / (src.stem + ".ptxas.log")).write_text(log)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{log[-6000:]}")
    return log


def _variant(out: Path, defines: list[str]) -> Path:
    srcs = [CSRC / n for n in SOURCES]
    headers = sorted(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "b200llm.h"]
    if not _stale(out, [*srcs, *headers]):
        return out
    flags = [f for f in FLAGS if f not in ("-Xptxas", "-v")]
    cmd = [NVCC, *flags, *defines, "-shared", "-o", str(out), *map(str, srcs), "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(r.stdout + r.stderr)
    return out


def build_trace_variant() -> Path:
    """Profiling-only library with the kernels' clock64 milestone traces compiled in."""
    return _variant(HERE / "libb200llm_trace.so", ["-DB200_ATTN_TRACE", "-DB200_STEP_TRACE", "-DB200_TUNE_KNOBS"])


def build_tune_variant() -> Path:
    """The product code plus the environment switches used for ablations and sweeps
    (B200LLM_SKIP, B200LLM_SPLITS, ...). The product library reads no
    environment variable; bench.py's decode_classes leg and profiles/*.sh load this one through
    B200LLM_LIB."""
    return _variant(HERE / "libb200llm_tune.so", ["-DB200_TUNE_KNOBS"])


def build(force: bool = False, verbose: bool = False) -> Path:
    BUILD.mkdir(exist_ok=True)
    headers = sorted(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "b200llm.h"]
    jobs = []
    for name in SOURCES:
        src = CSRC / name
        obj = BUILD / (src.stem + ".o")
        if force or _stale(obj, [src, *headers]):
            jobs.append((src, obj))
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as pool:
            for log in pool.map(lambda j: _compile(*j), jobs):
                if verbose:
                    print(log)
    objs = [BUILD / (Path(n).stem + ".o") for n in SOURCES]
    if jobs or not LIB.exists():
        cmd = [NVCC, "-shared", "-o", str(LIB), *map(str, objs), "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    if "--trace" in sys.argv:
        print(build_trace_variant())
        sys.exit(0)
    if "--tune" in sys.argv:
        print(build_tune_variant())
        sys.exit(0)
    lib = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(lib)
