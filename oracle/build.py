"""ORACLE — TEST INFRASTRUCTURE ONLY.  Build recipe for the checkers.

* ``build_oracle()``  -> oracle/_ref/libdibr_oracle.so : the C++ CPU restatement (dibr_oracle.cpp).
* ``build_ref_C()``   -> oracle/_ref/ref_C.so          : the reference's OWN CUDA implementation of the
  hot path (kaolin/csrc/render/mesh/{rasterization,dibr_soft_mask,deftet}{.cpp,_cuda.cu}), compiled
  unmodified for sm_100a from the sources where they lie under /root/reference, with the
  reference's nvcc flags (setup.py:216-220: -O3 -DWITH_CUDA).  Only possible in the build container
  (the GPU box has no /root/reference; it uses the prebuilt .so that travels with the snapshot).

Outputs go ONLY into oracle/_ref/ (git-ignored, not gpurun-ignored).
"""
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REFERENCE = os.environ.get("DIBR_REFERENCE_ROOT", "/root/reference")


def _run(cmd):
    print("[oracle/build]", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def build_oracle(force=False):
    os.makedirs(OUT, exist_ok=True)
    src = os.path.join(HERE, "dibr_oracle.cpp")
    out = os.path.join(OUT, "libdibr_oracle.so")
    if not force and _newer(out, [src]):
        return out
    # -ffp-contract=off: the only fused operations are the explicit std::fma calls.
    _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off",
          "-march=native", "-fvisibility=hidden", "-o", out, src])
    return out


def build_ref_C(force=False):
    """Returns the path of ref_C.so, or None when the reference tree is not present."""
    os.makedirs(OUT, exist_ok=True)
    out = os.path.join(OUT, "ref_C.so")
    srcdir = os.path.join(REFERENCE, "kaolin", "csrc", "render", "mesh")
    if not os.path.isdir(srcdir):
        return out if os.path.exists(out) else None
    import torch
    from torch.utils import cpp_extension
    cpp = [os.path.join(srcdir, "rasterization.cpp"), os.path.join(srcdir, "dibr_soft_mask.cpp"),
           os.path.join(srcdir, "deftet.cpp"), os.path.join(HERE, "ref_bindings.cpp")]
    cu = [os.path.join(srcdir, "rasterization_cuda.cu"), os.path.join(srcdir, "dibr_soft_mask_cuda.cu"),
          os.path.join(srcdir, "deftet_cuda.cu")]
    if not force and _newer(out, cpp + cu):
        return out
    tmp = os.path.join(OUT, "_obj")
    os.makedirs(tmp, exist_ok=True)
    inc = [f"-I{p}" for p in cpp_extension.include_paths(device_type="cuda")]
    inc += [f"-I{sysconfig.get_paths()['include']}", f"-I{REFERENCE}"]
    abi = f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"
    common = ["-O3", "-DWITH_CUDA", "-DTORCH_EXTENSION_NAME=ref_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
              abi, "-std=c++17"]
    objs = []
    for s in cpp:
        o = os.path.join(tmp, os.path.basename(s) + ".o")
        _run(["g++", "-fPIC", "-c", s, "-o", o] + common + inc)
        objs.append(o)
    for s in cu:
        o = os.path.join(tmp, os.path.basename(s) + ".o")
        _run(["nvcc", "-Xcompiler", "-fPIC", "-c", s, "-o", o, "-DTHRUST_IGNORE_CUB_VERSION_CHECK",
              "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo"] + common + inc)
        objs.append(o)
    libdirs = cpp_extension.library_paths(device_type="cuda")
    link = ["g++", "-shared", "-o", out] + objs
    for d in libdirs:
        link += [f"-L{d}", f"-Wl,-rpath,{d}"]
    link += ["-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart"]
    _run(link)
    shutil.rmtree(tmp, ignore_errors=True)
    return out


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_oracle(force))
    print(build_ref_C(force))
