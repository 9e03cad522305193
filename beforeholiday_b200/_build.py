"""In-tree build of the native pieces (no JIT cache: the .so files travel with the source tree).

  beforeholiday_b200/libmt_b200.so                      C-ABI CUDA library (nvcc, sm_100a, shared cudart)
  beforeholiday_b200/amp_C.cpython-*.so                 pybind11 adaptor over it (g++, torch headers)

`python -m beforeholiday_b200._build` builds both; __graft_entry__.build() calls build_all().
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import subprocess
import sys
import sysconfig
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
INCLUDE = PKG.parent / "include"
OBJ = PKG / "build"
NVCC = os.environ.get("MT_NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("MT_CXX", "/usr/bin/g++")

LIB_SOURCES = ["mt_table.cu", "mt_api_scale.cu", "mt_api_norm.cu", "mt_api_optim.cu", "mt_api_noalloc.cu", "mt_api_p2p.cu", "mt_api_more.cu", "mt_api_distlamb.cu"]
LIB_HEADERS = ["mt_common.cuh", "mt_ops.cuh", "mt_host.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    # Arithmetic contract with the oracle: IEEE div/sqrt, no FTZ, no fast-math.  (FMA contraction is blocked at the
    # source level by the __f*_rn intrinsics; -fmad=false makes the remaining plain expressions safe too.)
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
]


def lib_path() -> Path:
    return PKG / "libmt_b200.so"


def ext_path() -> Path:
    return PKG / ("amp_C" + sysconfig.get_config_var("EXT_SUFFIX"))


def _newer(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def _run(cmd, **kw):
    r = subprocess.run(cmd, capture_output=True, text=True, **kw)
    if r.returncode != 0:
        raise RuntimeError("build command failed:\n  " + " ".join(map(str, cmd)) + "\n" + r.stdout + r.stderr)
    return r


def build_lib(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(exist_ok=True)
    hdrs = [CSRC / h for h in LIB_HEADERS] + [INCLUDE / "mt_b200.h"]
    objs = []

    def compile_one(name: str) -> Path:
        src = CSRC / name
        obj = OBJ / (name + ".o")
        if force or _newer(obj, [src] + hdrs):
            cmd = [NVCC, "-c", str(src), "-o", str(obj)] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
            r = _run(cmd)
            if verbose:
                sys.stderr.write(r.stderr)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(len(LIB_SOURCES), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(compile_one, LIB_SOURCES))
    out = lib_path()
    if force or _newer(out, objs):
        _run([NVCC, "-shared", "-o", str(out)] + [str(o) for o in objs] +
             # the runtime is linked dynamically: inside a torch process the loader hands us the libcudart torch has
             # already mapped (one runtime instance for both), a bare C host finds the toolkit's through the rpath
             ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "shared", "-Xcompiler", "-fPIC",
              "-Xlinker", "-rpath", "-Xlinker", "/usr/local/cuda/lib64"])
    return out


def build_ext(force: bool = False) -> Path:
    import torch  # noqa: F401
    from torch.utils import cpp_extension as ce

    src = CSRC / "amp_C_module.cpp"
    out = ext_path()
    lib = build_lib()
    if not (force or _newer(out, [src, INCLUDE / "mt_b200.h", lib])):
        return out
    inc = ce.include_paths() + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]
    tlib = ce.library_paths()
    cmd = [CXX, "-O2", "-std=c++17", "-fPIC", "-shared", "-w", str(src), "-o", str(out),
           "-DTORCH_EXTENSION_NAME=amp_C", "-DTORCH_API_INCLUDE_EXTENSION_H"]
    cmd += [f"-I{p}" for p in inc]
    cmd += [f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    for p in tlib:
        cmd += [f"-L{p}", f"-Wl,-rpath,{p}"]
    cmd += [f"-L{PKG}", "-Wl,-rpath,$ORIGIN", "-lmt_b200",
            "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda"]
    _run(cmd)
    return out


def build_all(force: bool = False, verbose: bool = False):
    return build_lib(force, verbose), build_ext(force)


if __name__ == "__main__":
    force = "--force" in sys.argv
    verbose = "-v" in sys.argv
    for p in build_all(force, verbose):
        print("built", p)
