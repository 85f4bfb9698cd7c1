"""In-tree build of the native code.

Two artefacts, both written next to the Python sources so that they travel with the repo:

* ``_native/libb200dt.so``   : every ``csrc/*.cu`` compiled by plain ``nvcc`` for sm_100a and
  exposed through the C ABI of ``include/b200dt.h`` (bound with ctypes, see ``_lib.py``).
* ``_native/_reducer*.so``   : the C++ gradient reducer (``csrc/reducer/reducer.cpp``), a torch
  extension because it hooks into autograd.

``python -m pytorch_distributed_training_tutorials_b200.build`` builds both; nothing here needs a GPU.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
import sysconfig
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = ROOT / "csrc"
NATIVE = PKG / "_native"
BUILD = ROOT / "build"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC=/path/to/nvcc")


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def kernel_sources():
    return sorted(CSRC.glob("*.cu"))


def lib_path() -> Path:
    # B200DT_LIB: load another build of the kernel library (A/B measurements of kernel changes on one box)
    override = os.environ.get("B200DT_LIB")
    return Path(override) if override else NATIVE / "libb200dt.so"


def build_kernels(force: bool = False, verbose: bool = False) -> Path:
    """Compile csrc/*.cu -> _native/libb200dt.so (skipped when sources are unchanged)."""
    NATIVE.mkdir(exist_ok=True)
    BUILD.mkdir(exist_ok=True)
    srcs = kernel_sources()
    headers = sorted(CSRC.glob("*.cuh")) + sorted((ROOT / "include").glob("*.h"))
    stamp = NATIVE / "libb200dt.stamp"
    dig = _digest(srcs + headers)
    out = lib_path()
    if not force and out.exists() and stamp.exists() and stamp.read_text() == dig:
        return out
    nvcc = _nvcc()

    def compile_one(src: Path) -> Path:
        obj = BUILD / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(ROOT / "include"), "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    cuda_lib = Path(nvcc).resolve().parent.parent / "lib64"
    cmd = [nvcc, "-shared", "-o", str(out), *map(str, objs), "-cudart", "shared",
           "-L", str(cuda_lib), "-Xlinker", "-soname,libb200dt.so"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    return out


def reducer_path() -> Path | None:
    ext = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    p = NATIVE / f"_reducer{ext}"
    return p if p.exists() else None


def build_reducer(force: bool = False, verbose: bool = False) -> Path:
    """Compile csrc/reducer/reducer.cpp into a torch extension placed under _native/."""
    NATIVE.mkdir(exist_ok=True)
    srcs = sorted((CSRC / "reducer").glob("*.cpp"))
    if not srcs:
        raise RuntimeError("csrc/reducer/*.cpp not found")
    hdrs = sorted((CSRC / "reducer").glob("*.h")) + sorted((ROOT / "include").glob("*.h"))
    stamp = NATIVE / "_reducer.stamp"
    dig = _digest(srcs + hdrs)
    existing = reducer_path()
    if not force and existing is not None and stamp.exists() and stamp.read_text() == dig:
        return existing
    from torch.utils import cpp_extension

    build_dir = BUILD / "reducer"
    build_dir.mkdir(parents=True, exist_ok=True)
    cpp_extension.load(
        name="_reducer",
        sources=[str(s) for s in srcs],
        extra_include_paths=[str(ROOT / "include")],
        extra_cflags=["-O2", "-std=c++17"],
        with_cuda=True,                     # CUDA runtime headers + libcudart / c10_cuda (the reducer launches kernels)
        build_directory=str(build_dir),
        verbose=verbose,
        is_python_module=True,
    )
    built = next(build_dir.glob("_reducer*.so"))
    ext = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    dst = NATIVE / f"_reducer{ext}"
    shutil.copy2(built, dst)
    stamp.write_text(dig)
    return dst


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_kernels(force=force, verbose=verbose)
    if (CSRC / "reducer").exists():
        build_reducer(force=force, verbose=verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("built:", lib_path(), reducer_path())
