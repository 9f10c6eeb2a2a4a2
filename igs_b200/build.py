"""In-tree build of the native code (no JIT cache, so the .so files travel with the repo snapshot).

  igs_b200/lib/libigs_raster.so   hand-written sm_100a kernels + the C-ABI (include/igs_raster.h); no torch
  igs_b200/_C*.so                 thin pybind/torch glue linking libigs_raster.so (rpath $ORIGIN/lib)

Both are compiled with `-gencode arch=compute_100a,code=sm_100a -lineinfo` (nvcc cross-compiles
without a GPU).  `python -m igs_b200.build` or `__graft_entry__.build()`.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "lib" / "libigs_raster.so"
ROOT = HERE.parent


def ext_path() -> Path:
    suffix = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    return HERE / ("_C" + suffix)


def _newer(target: Path, sources) -> bool:
    if not target.exists():
        return False
    t = target.stat().st_mtime
    return all(Path(s).stat().st_mtime <= t for s in sources)


def build_lib(verbose: bool = False) -> Path:
    cmd = ["make", "-C", str(CSRC), "-j", str(os.cpu_count() or 4)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout[-4000:] + r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libigs_raster.so failed")
    return LIB


def build_ext(verbose: bool = False) -> Path:
    out = ext_path()
    src = CSRC / "torch_ext.cpp"
    hdr = ROOT / "include" / "igs_raster.h"
    if _newer(out, [src, hdr, LIB]):
        return out
    import torch
    from torch.utils import cpp_extension as ce

    inc = ce.include_paths(device_type="cuda")
    libdirs = ce.library_paths(device_type="cuda")
    pyinc = sysconfig.get_paths()["include"]
    cxx = os.environ.get("CXX", "g++")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-Wno-deprecated-declarations"]
    cmd += [f"-I{p}" for p in inc] + [f"-I{pyinc}", "-I/usr/local/cuda/include"]
    cmd += [str(src), "-o", str(out)]
    cmd += [f"-L{p}" for p in libdirs] + [f"-L{LIB.parent}", "-L/usr/local/cuda/lib64"]
    cmd += ["-ligs_raster", "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
    cmd += ["-Wl,-rpath,$ORIGIN/lib"] + [f"-Wl,-rpath,{p}" for p in libdirs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout[-4000:] + r.stderr[-6000:])
    if r.returncode != 0:
        raise RuntimeError("building igs_b200._C failed")
    return out


def build_all(verbose: bool = False) -> None:
    build_lib(verbose)
    build_ext(verbose)


if __name__ == "__main__":
    build_all(verbose="-v" in sys.argv)
    print(LIB, ext_path())
