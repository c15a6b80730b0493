"""In-tree native build: nvcc (sm_100a) for the CUDA kernel library, g++ for the C++ control-plane core.

Artifacts land in ``determined_b200/_lib/`` (git-ignored, but shipped to the GPU box by gpurun):

* ``libdet_kernels.so`` — every hand-written CUDA kernel behind a C ABI (``include/det_kernels.h``)
* ``libdet_core.so``    — C++ core (nprand, searchers, sequencer, schedulers, config) behind a C ABI
* ``det-master`` / ``det-agent`` — the single-box master daemon and node agent executables

The build is incremental (object files keyed by source mtime) and needs no GPU: nvcc cross-compiles.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys
from pathlib import Path
from typing import Optional,  List, Sequence

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIB_DIR = Path(__file__).resolve().parent / "_lib"
OBJ_DIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-I", str(ROOT / "include"),
    "-I", str(CSRC / "kernels"),
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread", "-I", str(ROOT / "include"), "-I", str(CSRC / "core")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _newer(src: Path, deps: Sequence[Path], out: Path) -> bool:
    if not out.exists():
        return True
    t = out.stat().st_mtime
    return any(d.stat().st_mtime > t for d in [src, *deps])


def _run(cmd: List[str]) -> None:
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("build command failed:\n  " + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)


def cuda_sources() -> List[Path]:
    out: List[Path] = []
    for sub in ("kernels", "optim", "comm"):
        out += sorted((CSRC / sub).glob("*.cu"))
    return out


def core_sources() -> List[Path]:
    return sorted((CSRC / "core").glob("*.cpp"))


def build_kernels(force: bool = False, verbose: bool = False) -> Path:
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    headers = list((CSRC / "kernels").glob("*.cuh")) + list((CSRC / "kernels").glob("*.h")) + list((ROOT / "include").glob("*.h"))
    srcs = cuda_sources()
    objs = []
    jobs = []
    for s in srcs:
        o = OBJ_DIR / (s.parent.name + "_" + s.stem + ".o")
        objs.append(o)
        if force or _newer(s, headers, o):
            jobs.append([nvcc, *NVCC_FLAGS, "-c", str(s), "-o", str(o)])
    if verbose and jobs:
        print(f"[build] compiling {len(jobs)} CUDA translation units", file=sys.stderr)
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(_run, jobs))
    lib = LIB_DIR / "libdet_kernels.so"
    if force or jobs or not lib.exists():
        _run([nvcc, "-shared", "-o", str(lib), *map(str, objs), "-lcudart"])
    return lib


def build_core(force: bool = False, verbose: bool = False) -> Path:
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX", "g++")
    headers = list((CSRC / "core").glob("*.h")) + list((CSRC / "core").glob("*.hpp"))
    srcs = core_sources()
    lib = LIB_DIR / "libdet_core.so"
    if not srcs:
        return lib
    objs, jobs = [], []
    for s in srcs:
        o = OBJ_DIR / ("core_" + s.stem + ".o")
        objs.append(o)
        if force or _newer(s, headers, o):
            jobs.append([cxx, *CXX_FLAGS, "-c", str(s), "-o", str(o)])
    if verbose and jobs:
        print(f"[build] compiling {len(jobs)} C++ translation units", file=sys.stderr)
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(_run, jobs))
    lib_objs = [o for o, s in zip(objs, srcs) if not s.stem.startswith("main_")]
    if force or jobs or not lib.exists():
        _run([cxx, "-shared", "-o", str(lib), *map(str, lib_objs), "-pthread", "-lssl", "-lcrypto"])
    for o, s in zip(objs, srcs):
        if s.stem.startswith("main_"):
            exe = LIB_DIR / ("det-" + s.stem[len("main_"):].replace("_", "-"))
            if force or jobs or not exe.exists():
                _run([cxx, "-o", str(exe), str(o), *map(str, lib_objs), "-pthread", "-lssl", "-lcrypto"])
    return lib


def build_selftest_tsan(verbose: bool = False) -> Path:
    """``det-selftest`` compiled together with the whole control-plane library under ThreadSanitizer
    (``build/det-selftest-tsan``). Run it directly; it exits non-zero on a failed expectation or a reported race."""
    out = ROOT / "build" / "det-selftest-tsan"
    out.parent.mkdir(parents=True, exist_ok=True)
    srcs = [s for s in core_sources() if not s.stem.startswith("main_") or s.stem == "main_selftest"]
    flags = [f for f in CXX_FLAGS if f not in ("-O2", "-O3", "-fPIC")]
    last: Optional[Exception] = None
    # not every toolchain ships the sanitizer runtimes: try $CXX first, then the distribution compiler
    for cxx in dict.fromkeys([os.environ.get("CXX") or "g++", "g++", "/usr/bin/g++"]):
        cmd = [cxx, *flags, "-O1", "-g", "-fsanitize=thread", "-fno-omit-frame-pointer", *map(str, srcs), "-o", str(out), "-pthread", "-lssl", "-lcrypto"]
        if verbose:
            print("[build] " + " ".join(cmd[:12]) + " ...", file=sys.stderr)
        try:
            _run(cmd)
            return out
        except (RuntimeError, FileNotFoundError) as e:
            last = e
    raise RuntimeError(f"no compiler with a ThreadSanitizer (tsan) runtime found: {last}")


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_kernels(force=force, verbose=verbose)
    build_core(force=force, verbose=verbose)


if __name__ == "__main__":
    if "--tsan" in sys.argv:
        print("built:", build_selftest_tsan(verbose=True))
        sys.exit(0)
    build_all(force="--force" in sys.argv, verbose=True)
    print("built:", sorted(p.name for p in LIB_DIR.iterdir()))
