"""In-tree build of the native pieces (nvcc for sm_100a; gcc/g++ for the test-side libraries).

  libdrt.so                        the product: C ABI (include/drt.h) + CUDA kernels     [nvcc, sm_100a]
  oracle/_build/libdrt_oracle.so   CPU restatement of the reference path (checker only)   [gcc]
  oracle/_ref/libdrt_ref.so        the reference's own sources compiled host-side         [g++; needs /root/reference]
  tests/hostemu/_build/...         device headers compiled as plain C++ for logic tests   [g++]

The built .so files stay in-tree (git-ignored) so they travel to the GPU box with the snapshot.
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libdrt.so")
HOSTEMU = os.path.join(ROOT, "tests", "hostemu", "_build", "libdrt_hostemu.so")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libdrt_ref.so")
PORT_LIB = os.path.join(ROOT, "oracle", "_build", "libdrt_oracle.so")

NVCC_FLAGS = [
    "-std=c++17", "-O3", "-lineinfo",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC,-fno-fast-math,-ffp-contract=off",
]
NVCC_FLAGS += os.environ.get("DRT_NVCC_EXTRA", "").split()  # measurement builds only (e.g. -DDRT_RG_CTX=384)
CUDA_SOURCES = ["drt_kernels.cu", "drt_api.cu"]
HOST_SOURCES = ["drt_gltf.cpp", "drt_gltf_robust.cpp", "drt_png.cpp", "drt_jpeg.cpp"]


def _host_cxx():
    # the image exports CXX=/opt/gcc/bin/g++ (a wrapper without libgomp.spec); prefer the distro compiler
    return "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _all_csrc():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "drt.h")]


def _run(cmd, cwd=None):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build command failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return r.stdout


def build_libdrt(force=False, verbose=False):
    """nvcc -> dustraytracer_b200/libdrt.so (sm_100a only)."""
    if not force and not _newer(LIB, _all_csrc()):
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    objdir = os.path.join(PKG, "build")
    os.makedirs(objdir, exist_ok=True)
    objs = []
    for src in CUDA_SOURCES + HOST_SOURCES:
        obj = os.path.join(objdir, src + ".o")
        extra = ["-Xptxas", "-v"] if (verbose and src.endswith(".cu")) else []
        out = _run([nvcc] + NVCC_FLAGS + extra + ["-ccbin", _host_cxx(), "-c", os.path.join(CSRC, src), "-o", obj])
        if verbose and out:
            print(out)
        objs.append(obj)
    _run([nvcc, "-shared", "-ccbin", _host_cxx(), "-Xlinker", "-Bsymbolic", "-o", LIB] + objs + ["-lz"])
    return LIB


def build_hostemu(force=False):
    src = os.path.join(ROOT, "tests", "hostemu", "hostemu.cpp")
    if not force and not _newer(HOSTEMU, _all_csrc() + [src]):
        return HOSTEMU
    os.makedirs(os.path.dirname(HOSTEMU), exist_ok=True)
    _run([_host_cxx(), "-std=c++17", "-O2", "-fPIC", "-fopenmp", "-ffp-contract=off", "-x", "c++", "-shared", "-Wl,-Bsymbolic", "-o", HOSTEMU,
          src] + [os.path.join(CSRC, f) for f in HOST_SOURCES] + ["-lz"])
    return HOSTEMU


def build_oracle(force=False):
    """Compiles the checker libraries. The reference build needs /root/reference (build container only)."""
    odir = os.path.join(ROOT, "oracle")
    built = []
    if os.path.exists(os.path.join(odir, "drt_oracle.c")):
        _run(["make", "-C", odir, "port"] + (["-B"] if force else []))
        built.append(PORT_LIB)
    if os.path.isdir("/root/reference/DustRayTracer/src"):
        _run(["make", "-C", odir, "ref"] + (["-B"] if force else []))
        built.append(REF_LIB)
        # the reference's CUDA path for the GPU box (measurement baseline, ~3 min the first time)
        _run(["make", "-C", odir, "refcuda"] + (["-B"] if force else []))
        built += [os.path.join(odir, "_ref", "libdrt_refcuda.so"), os.path.join(odir, "_ref", "libdrt_refcuda_fast.so")]
    return built


CLI = os.path.join(PKG, "drt_render")


def build_cli(force=False):
    """g++ -> dustraytracer_b200/drt_render: command-line front end over include/drt.hpp (links libdrt.so by rpath)."""
    src = os.path.join(ROOT, "tools", "drt_render.cpp")
    deps = [src, os.path.join(ROOT, "include", "drt.hpp"), os.path.join(ROOT, "include", "drt.h"), LIB]
    if not force and not _newer(CLI, deps):
        return CLI
    _run([_host_cxx(), "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"), src, "-o", CLI, "-L", PKG, "-ldrt", "-Wl,-rpath,$ORIGIN"])
    return CLI


def build_all(force=False, verbose=False):
    out = [build_libdrt(force, verbose), build_hostemu(force), build_cli(force)]
    out += build_oracle(force)
    return out


if __name__ == "__main__":
    for p in build_all(force="--force" in sys.argv, verbose="-v" in sys.argv):
        print("built", os.path.relpath(p, ROOT))
