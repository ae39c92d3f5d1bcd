"""Builds k3m_b200/libk3m_sm100a.so from k3m_b200/csrc/*.cu with nvcc (sm_100a only, in-tree)."""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libk3m_sm100a.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC"] + os.environ.get("K3M_NVCC_EXTRA", "").split()    # e.g. -DK3M_GEMM_PROBE_BUILD


def sources():
    return sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))


STAMP = os.path.join(HERE, "csrc", "build", "flags.txt")


def needs_build():
    if not os.path.exists(LIB):
        return True
    # a library built with other flags (e.g. a -DK3M_GEMM_PROBE_BUILD measurement build) is stale for this run
    if not os.path.exists(STAMP) or open(STAMP).read() != " ".join(FLAGS):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(HERE, "csrc", "*.cuh"))
    return any(os.path.getmtime(s) > t for s in deps)


# The fp32 CHECK library (tests only): the same row-wise / embedding / loss kernel sources compiled with fp32 storage
# (-DK3M_ACT_F32) plus CUDA-core fp32 GEMM / attention behind the same entry points (gemm_f32.cu, attention_f32.cu);
# the tcgen05 GEMM and the mma.sync attention are bf16 by construction and stay out of it.
LIB_F32 = os.path.join(HERE, "libk3m_sm100a_f32.so")
F32_EXCLUDE = ("gemm_sm100.cu", "attention.cu")


def _compile(srcs, objdir, flags, lib, verbose):
    os.makedirs(objdir, exist_ok=True)
    procs, objs = [], []
    for src in srcs:
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [NVCC] + flags + ["-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd))
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out.decode())
            raise RuntimeError("nvcc failed on %s" % src)
    subprocess.check_call([NVCC, "-shared", "-o", lib] + objs + ["-lcudart"])


def build(force=False, verbose=False):
    if not force and not needs_build() and os.path.exists(LIB_F32) and os.path.getmtime(LIB_F32) >= os.path.getmtime(LIB):
        return LIB
    _compile(sources(), os.path.join(HERE, "csrc", "build"), FLAGS, LIB, verbose)
    _compile([s for s in sources() if os.path.basename(s) not in F32_EXCLUDE], os.path.join(HERE, "csrc", "build", "f32"),
             FLAGS + ["-DK3M_ACT_F32"], LIB_F32, verbose)
    with open(STAMP, "w") as f:
        f.write(" ".join(FLAGS))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
