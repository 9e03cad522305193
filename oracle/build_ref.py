#!/usr/bin/env python
"""Build the REFERENCE's own multi-tensor extensions, unmodified, into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing under beforeholiday_b200/ imports this or the
libraries it produces; only tests/, __graft_entry__.smoke(), bench.py's
reference legs and the measurement scripts under tools/ (bench_ops.py: the same baseline legs for the other configs)
may load oracle/_ref/*.so, and only as the checker / baseline.

What is built (sources are compiled where they lie under /root/reference; none
are copied into this repository):

  oracle/_ref/amp_C<EXT_SUFFIX>                 <- csrc/amp_C_frontend.cpp + csrc/multi_tensor_*.cu
                                                   (reference recipe: setup.py:313-336, flags :310-311)
  oracle/_ref/distributed_adam_cuda<EXT_SUFFIX> <- apex/contrib/csrc/optimizers/multi_tensor_distopt_adam{.cpp,_kernel.cu}
                                                   (reference recipe: setup.py, "--distributed_adam")
  oracle/_ref/distributed_lamb_cuda<EXT_SUFFIX> <- apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb{.cpp,_kernel.cu}
                                                   (reference recipe: setup.py, "--distributed_lamb")

The reference's setup.py cannot run on a CUDA build of torch (setup.py:157 parses
torch.version.hip unconditionally), so the same flags are issued directly with
nvcc/g++ for sm_100a.  /root/reference does not exist on the GPU box: this script
only runs in the build container; the resulting .so files travel with the snapshot
(oracle/_ref is git-ignored but not gpurun-ignored).
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import subprocess
import sys
import sysconfig
from pathlib import Path

REF = Path(os.environ.get("MT_REFERENCE_ROOT", "/root/reference"))
HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"
OBJ = OUT / "obj"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")


def _torch_paths():
    import torch  # noqa: F401
    from torch.utils import cpp_extension as ce

    inc = ce.include_paths() + [sysconfig.get_paths()["include"]]
    lib = ce.library_paths()
    return inc, lib


def _common_defs(name: str):
    return [
        f"-DTORCH_EXTENSION_NAME={name}",
        "-DTORCH_API_INCLUDE_EXTENSION_H",
        "-DVERSION_GE_1_1",
        "-DVERSION_GE_1_3",
        "-DVERSION_GE_1_5",
    ]


def _compile(src: Path, name: str, inc, extra_inc):
    OBJ.mkdir(parents=True, exist_ok=True)
    obj = OBJ / (name + "__" + src.name + ".o")
    if obj.exists() and obj.stat().st_mtime >= src.stat().st_mtime:
        return obj
    incs = [f"-I{p}" for p in list(inc) + list(extra_inc)]
    if src.suffix == ".cu":
        cmd = [
            NVCC, "-c", str(src), "-o", str(obj),
            "-gencode", "arch=compute_100a,code=sm_100a",
            "-O3", "--use_fast_math", "-lineinfo", "-std=c++17",
            "--expt-relaxed-constexpr", "--expt-extended-lambda",
            "-Xcompiler", "-fPIC", "-w",
            # what torch.utils.cpp_extension.CUDAExtension always adds (COMMON_NVCC_FLAGS); without them
            # c10::Half comparisons in multi_tensor_distopt_lamb_kernel.cu:352 are ambiguous
            "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
            "-D__CUDA_NO_BFLOAT16_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__",
        ] + _common_defs(name) + incs
    else:
        cmd = [CXX, "-c", str(src), "-o", str(obj), "-O3", "-std=c++17", "-fPIC", "-w"] \
            + _common_defs(name) + incs
    subprocess.run(cmd, check=True)
    return obj


def _link(objs, name: str, lib):
    out = OUT / (name + sysconfig.get_config_var("EXT_SUFFIX"))
    cmd = [CXX, "-shared", "-o", str(out)] + [str(o) for o in objs]
    for p in lib:
        cmd += [f"-L{p}", f"-Wl,-rpath,{p}"]
    cmd += ["-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64",
            "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
    subprocess.run(cmd, check=True)
    return out


def build(jobs: int | None = None) -> list[Path]:
    if not REF.exists():
        raise SystemExit(f"{REF} not present: oracle/_ref can only be built in the build container")
    inc, lib = _torch_paths()
    csrc = REF / "csrc"
    targets = {
        "amp_C": [csrc / "amp_C_frontend.cpp"] + sorted(csrc.glob("multi_tensor_*.cu")),
        "distributed_adam_cuda": [
            REF / "apex/contrib/csrc/optimizers/multi_tensor_distopt_adam.cpp",
            REF / "apex/contrib/csrc/optimizers/multi_tensor_distopt_adam_kernel.cu",
        ],
        "distributed_lamb_cuda": [
            REF / "apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp",
            REF / "apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu",
        ],
    }
    jobs = jobs or os.cpu_count() or 4
    outs = []
    with cf.ThreadPoolExecutor(max_workers=jobs) as ex:
        futs = {name: [ex.submit(_compile, s, name, inc, [csrc]) for s in srcs] for name, srcs in targets.items()}
        for name, fl in futs.items():
            objs = [f.result() for f in fl]
            outs.append(_link(objs, name, lib))
    return outs


if __name__ == "__main__":
    for o in build():
        print("built", o)
    sys.exit(0)
