"""Builds cariboulite_b200/libcaribou_cuda.so in-tree with nvcc for sm_100a (no GPU needed)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# CBL_VARIANT=<tag> (+ CBL_NVCC_EXTRA="-D...") builds libcaribou_cuda_<tag>.so beside the product library
VARIANT = os.environ.get("CBL_VARIANT", "")
OBJ = os.path.join(HERE, "_obj" + ("_" + VARIANT if VARIANT else ""))
LIB = os.path.join(HERE, "libcaribou_cuda%s.so" % ("_" + VARIANT if VARIANT else ""))
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CU = ["unpack.cu", "iir.cu", "iir_k1.cu", "iir_k2.cu", "iir_k3.cu", "iir_k4.cu", "iir_k1y.cu", "iir_k2y.cu", "iir_k3y.cu",
      "iir_k4y.cu", "iir_r1.cu", "iir_r2.cu", "iir_r3.cu", "iir_r4.cu", "iir_ra1.cu", "iir_ra2.cu", "iir_ra3.cu", "iir_ra4.cu", "tx.cu", "api.cu", "shard.cu", "compat.cu", "noalloc.cu"]
CPP = ["tables.cpp", "ring.cpp"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall"]


def _newer(src, dst):
    if not os.path.exists(dst):
        return True
    t = os.path.getmtime(dst)
    deps = [src] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(HERE, "..", "include", "caribou_cuda.h"))
    deps.append(os.path.abspath(__file__))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _compile(name):
    src = os.path.join(CSRC, name)
    obj = os.path.join(OBJ, name + ".o")
    if not _newer(src, obj):
        return obj
    extra = os.environ.get("CBL_NVCC_EXTRA", "").split()
    cmd = [NVCC] + ARCH + FLAGS + extra + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (name, r.stdout, r.stderr))
    return obj


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    srcs = [f for f in CU + CPP if os.path.exists(os.path.join(CSRC, f))]
    if force:
        for f in os.listdir(OBJ):
            os.remove(os.path.join(OBJ, f))
    with ThreadPoolExecutor(max_workers=min(12, len(srcs))) as ex:
        objs = list(ex.map(_compile, srcs))
    if force or not os.path.exists(LIB) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-lcudart_static", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    if verbose:
        print("built", LIB)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
