"""In-tree native build: nvcc -> lib/libdtf_b200_cuda.so (sm_100a), g++ -> lib/libdtf_b200_io.so.

The reference has no native code; these libraries replace what it delegated to the TF1
runtime (cuDNN/cuBLAS kernels, gRPC push/pull, Saver, TFRecord reader).
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIBDIR = os.path.join(PKG, "lib")
CUDA_LIB = os.path.join(LIBDIR, "libdtf_b200_cuda.so")
IO_LIB = os.path.join(LIBDIR, "libdtf_b200_io.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--use_fast_math", "-cudart", "static",
]


def _newer(srcs, out):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(s) > t for s in srcs)


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def build_cuda(force=False, verbose=False):
    srcs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    if not force and not _newer(deps, CUDA_LIB):
        return CUDA_LIB
    os.makedirs(LIBDIR, exist_ok=True)
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(LIBDIR, os.path.basename(s) + ".o")
        objs.append(o)
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), out))
    cmd = [_nvcc(), "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a", "-o", CUDA_LIB] + objs + ["-lrt", "-lpthread", "-ldl"]
    subprocess.run(cmd, check=True)
    for o in objs:
        os.remove(o)
    return CUDA_LIB


def build_io(force=False):
    iosrc = os.path.join(PKG, "io", "csrc")
    srcs = sorted(os.path.join(iosrc, f) for f in os.listdir(iosrc) if f.endswith(".cc"))
    deps = srcs + [os.path.join(iosrc, f) for f in os.listdir(iosrc) if f.endswith(".h")]
    if not force and not _newer(deps, IO_LIB):
        return IO_LIB
    os.makedirs(LIBDIR, exist_ok=True)
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", IO_LIB] + srcs
    subprocess.run(cmd, check=True)
    return IO_LIB


def build_io_sanitized():
    """AddressSanitizer + UndefinedBehaviorSanitizer build of the I/O library (lib/libdtf_b200_io_asan.so); run the
    I/O tests against it with scripts/asan_io.sh (the host-side sanitizer tier)."""
    iosrc = os.path.join(PKG, "io", "csrc")
    srcs = sorted(os.path.join(iosrc, f) for f in os.listdir(iosrc) if f.endswith(".cc"))
    out = os.path.join(LIBDIR, "libdtf_b200_io_asan.so")
    os.makedirs(LIBDIR, exist_ok=True)
    cmd = ["g++", "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-pthread", "-fno-omit-frame-pointer",
           "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-o", out] + srcs
    subprocess.run(cmd, check=True)
    return out


def build_all(force=False, verbose=False):
    return build_cuda(force, verbose), build_io(force)


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
