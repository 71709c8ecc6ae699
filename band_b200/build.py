"""In-tree build of the native libraries (no JIT cache: the .so files travel with the repo snapshot).

  band_b200/lib/libbandcu.so   CUDA shim, sm_100a only          (include/bandcu.h)
  band_b200/lib/libband_c.so   C++ host layer + Band C API      (include/band_c_api.h), links libbandcu.so
  oracle/_build/liborc.so      CPU oracle (test infrastructure) (oracle/Makefile)

nvcc cross-compiles for sm_100a without a GPU. A target is rebuilt only when one
of its sources is newer than the library.
"""
import glob
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "band_b200", "csrc")
LIBDIR = os.path.join(ROOT, "band_b200", "lib")
INCLUDE = os.path.join(ROOT, "include")

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, cwd=None):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return r.stdout


def build_bandcu(force=False, verbose=False):
    """libbandcu.so: every .cu under csrc/cuda, one nvcc invocation, sm_100a, -lineinfo."""
    out = os.path.join(LIBDIR, "libbandcu.so")
    cu = sorted(glob.glob(os.path.join(CSRC, "cuda", "*.cu")))
    deps = cu + glob.glob(os.path.join(CSRC, "cuda", "*.cuh")) + glob.glob(os.path.join(CSRC, "cuda", "*.h")) + \
        [os.path.join(INCLUDE, "bandcu.h")]
    if not force and not _stale(out, deps):
        return out
    os.makedirs(LIBDIR, exist_ok=True)
    cmd = [_nvcc()] + NVCC_ARCH + ["-lineinfo", "-O3", "-std=c++17", "-shared", "-Xcompiler",
                                   "-fPIC,-fvisibility=hidden", "-I" + INCLUDE, "-I" + os.path.join(CSRC, "cuda")]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += ["-o", out] + cu
    log = _run(cmd)
    if verbose:
        print(log)
    return out


def build_band_c(force=False):
    """libband_c.so: the C++ host layer (band::Buffer, operators, processor, C API) over libbandcu.so."""
    out = os.path.join(LIBDIR, "libband_c.so")
    srcs = sorted(glob.glob(os.path.join(CSRC, "band", "**", "*.cc"), recursive=True))
    if not srcs:
        return None
    hdrs = glob.glob(os.path.join(CSRC, "**", "*.h"), recursive=True) + glob.glob(os.path.join(INCLUDE, "*.h"))
    cu_lib = os.path.join(LIBDIR, "libbandcu.so")
    if not force and not _stale(out, srcs + hdrs + [cu_lib]):
        return out
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-fno-exceptions", "-Wall", "-Wextra",
           "-Wno-unused-parameter", "-I" + CSRC, "-I" + os.path.join(CSRC, "compat"),
           "-I" + INCLUDE, "-o", out] + srcs + ["-L" + LIBDIR, "-lbandcu", "-Wl,-rpath,$ORIGIN", "-lpthread"]
    _run(cmd)
    return out


def build_testing(force=False):
    """lib/libband_testing.so: C entry points that replay the reference tests' C++ call sequences
    (tests/cc/band_testing.cc) against libband_c.so. Test support only."""
    out = os.path.join(LIBDIR, "libband_testing.so")
    src = os.path.join(ROOT, "tests", "cc", "band_testing.cc")
    if not os.path.exists(src):
        return None
    host = os.path.join(LIBDIR, "libband_c.so")
    if not force and not _stale(out, [src, host]):
        return out
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    _run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-fno-exceptions", "-Wall", "-I" + CSRC,
          "-I" + os.path.join(CSRC, "compat"), "-I" + INCLUDE, "-o", out, src, "-L" + LIBDIR, "-lband_c", "-lbandcu",
          "-Wl,-rpath,$ORIGIN"])
    return out


def build_tools(force=False):
    """lib/band_preproc_benchmark: the C++ benchmark CLI (tools/band_preproc_benchmark.cc), a client of the two C ABIs."""
    out = os.path.join(LIBDIR, "band_preproc_benchmark")
    src = os.path.join(ROOT, "tools", "band_preproc_benchmark.cc")
    if not os.path.exists(src):
        return None
    host = os.path.join(LIBDIR, "libband_c.so")
    if not force and not _stale(out, [src, host]):
        return out
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    _run([cxx, "-O2", "-std=c++17", "-Wall", "-Wextra", "-I" + INCLUDE, "-o", out, src, "-L" + LIBDIR, "-lband_c", "-lbandcu",
          "-Wl,-rpath,$ORIGIN", "-lpthread"])
    return out


def build_oracle(force=False):
    """oracle/_build/liborc.so (checker + timed CPU baseline; never loaded by the product)."""
    odir = os.path.join(ROOT, "oracle")
    if force:
        _run(["make", "-C", odir, "clean"])
    arch = "x86-64-v3"
    try:
        with open("/proc/cpuinfo") as f:
            if " avx2" not in f.read():
                arch = "x86-64"
    except OSError:
        pass
    _run(["make", "-C", odir, "ORC_ARCH=" + arch])
    return os.path.join(odir, "_build", "liborc.so")


def build_all(force=False, verbose=False):
    libs = [build_bandcu(force, verbose)]
    b = build_band_c(force)
    if b:
        libs.append(b)
    t = build_testing(force)
    if t:
        libs.append(t)
    t = build_tools(force)
    if t:
        libs.append(t)
    libs.append(build_oracle(force))
    return libs


if __name__ == "__main__":
    import sys
    for p in build_all(force="--force" in sys.argv, verbose="-v" in sys.argv):
        print(p)
