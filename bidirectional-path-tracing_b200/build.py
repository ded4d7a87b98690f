"""In-tree build of libtrb.so (host C++ + sm_100a CUDA kernels + the C ABI of include/trb.h).

    python bidirectional-path-tracing_b200/build.py [--force] [--counters]

nvcc cross-compiles for sm_100a without a GPU.  The library is written next to this file so that it
travels to the GPU box with the gpurun snapshot (it is git-ignored, not gpurun-ignored).
Flags that matter for parity (DESIGN.md "Numerics"): -fmad=false on the device side and
-ffp-contract=off on the host side keep fp32 arithmetic un-contracted like the reference's x86-64 build.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
ROOT = PKG.parent
OBJ = PKG / "build"
LIB = PKG / "libtrb.so"
LIB_COUNTERS = PKG / "libtrb_counters.so"  # same sources with -DTRB_COUNTERS (node / triangle visit counts)

HOST_SRCS = ["host/obj_loader.cpp", "host/scene_build.cpp", "host/camera.cpp", "host/toml_exr.cpp", "host/checkpoint_png.cpp", "host/host_abi.cpp",
             "host/tinyrender_api.cpp"]
CUDA_SRCS = ["trb_abi.cu"]
CLI_SRC = "host/tinyrender_main.cpp"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off", "--expt-relaxed-constexpr", "-Xptxas", "-v",
]
CXX_FLAGS = ["-std=c++17", "-O2", "-fPIC", "-ffp-contract=off", "-Wall"]


def _stdcxx_link_args() -> list[str]:
    # this image's g++ only finds a static libstdc++.a; link the system DSO so that the library can live in
    # the same process as torch/numpy
    for p in ("/usr/lib/x86_64-linux-gnu/libstdc++.so.6", "/usr/lib64/libstdc++.so.6"):
        if os.path.exists(p):
            return ["-nostdlib++", p]
    return []


def _nvcc() -> str:
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: the CUDA extension cannot be built (there is no CPU fallback)")


def _run(cmd: list[str], log: Path | None = None) -> None:
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.write_text(r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build step failed: " + " ".join(cmd))


def _digest(paths: list[Path], extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(str(p.relative_to(ROOT)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def build(force: bool = False, counters: bool = False, verbose: bool = False, variant: str = "", defines: tuple = ()) -> Path:
    """Compile everything that is out of date; returns the path of libtrb.so.
    variant/defines: tuning experiments - the same sources with extra -D flags into libtrb_<variant>.so (use with TRB_LIB=...)."""
    OBJ.mkdir(exist_ok=True)
    lib_out = LIB_COUNTERS if counters else LIB
    sfx = "_cnt" if counters else ""
    if variant:
        lib_out, sfx = PKG / f"libtrb_{variant}.so", "_" + variant
    deps = [p for p in CSRC.rglob("*") if p.suffix in (".cu", ".cuh", ".cpp", ".h")] + [ROOT / "include" / "trb.h"]
    tag = _digest(deps, ("counters" if counters else "plain") + variant + " ".join(defines))
    stamp = OBJ / f"stamp{sfx}.txt"
    if not force and lib_out.exists() and stamp.exists() and stamp.read_text() == tag:
        return lib_out
    objs = []
    for src in HOST_SRCS:
        o = OBJ / (Path(src).stem + ".o")
        _run(["g++", *CXX_FLAGS, "-I", str(ROOT / "include"), "-c", str(CSRC / src), "-o", str(o)])
        objs.append(str(o))
    nvcc = _nvcc()
    for src in CUDA_SRCS:
        if not (CSRC / src).exists():
            continue
        o = OBJ / (Path(src).stem + sfx + ".o")
        flags = list(NVCC_FLAGS) + (["-DTRB_COUNTERS=1"] if counters else []) + list(defines)
        _run([nvcc, *flags, "-I", str(ROOT / "include"), "-c", str(CSRC / src), "-o", str(o)], log=OBJ / (Path(src).stem + sfx + ".ptxas.log"))
        objs.append(str(o))
    have_cuda = any((CSRC / s).exists() for s in CUDA_SRCS)
    link = ["g++", "-shared", *objs, "-o", str(lib_out), *_stdcxx_link_args(), "-lm", "-lpthread"]
    if have_cuda:
        cuda_lib = Path(nvcc).resolve().parent.parent / "lib64"
        link += ["-L", str(cuda_lib), "-lcudart_static", "-ldl", "-lrt"]
    _run(link)
    if (CSRC / CLI_SRC).exists() and not counters and not variant:
        _run(["g++", *CXX_FLAGS, "-I", str(ROOT / "include"), str(CSRC / CLI_SRC), "-o", str(PKG / "tinyrender_b200"),
              "-L", str(PKG), "-l:libtrb.so", "-Wl,-rpath,$ORIGIN", *_stdcxx_link_args(), "-lm"])
    stamp.write_text(tag)
    if verbose:
        print(f"built {lib_out}")
    return lib_out


if __name__ == "__main__":
    if "--variant" in sys.argv:  # python build.py --variant t4 -DTRB_FETCH_T=4 ...
        i = sys.argv.index("--variant")
        build(variant=sys.argv[i + 1], defines=tuple(a for a in sys.argv[i + 2:] if a.startswith("-D")), verbose=True)
    else:
        build(force="--force" in sys.argv, counters="--counters" in sys.argv, verbose=True)
