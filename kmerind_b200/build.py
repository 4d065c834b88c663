"""Builds libkmi.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

One object per supported k-mer length (kmi_ops_inst.cu with -DKMI_K=...) plus the API object,
compiled in parallel, linked into kmerind_b200/libkmi.so. No GPU is needed to build.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# KMI_VARIANT=name (tuning experiments only): separate object directory and libkmi_name.so, picked up by KMI_LIB
VARIANT = os.environ.get("KMI_VARIANT", "")
OUT = os.path.join(HERE, "libkmi_%s.so" % VARIANT if VARIANT else "libkmi.so")
KS = (15, 21, 31, 35, 47, 63)
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
         "-Xcompiler", "-fPIC", "-diag-suppress", "62,63"] + os.environ.get("KMI_NVCC_FLAGS", "").split()


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


def build(force=False, verbose=False):
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".hpp"))]
    hdrs.append(os.path.join(HERE, "..", "include", "kmi.h"))
    objdir = os.path.join(CSRC, "build", VARIANT) if VARIANT else os.path.join(CSRC, "build")
    os.makedirs(objdir, exist_ok=True)
    jobs = []
    objs = []
    inst = os.path.join(CSRC, "kmi_ops_inst.cu")
    for k in KS:
        o = os.path.join(objdir, "kmi_ops_k%d.o" % k)
        objs.append(o)
        if force or _newer(o, hdrs + [inst]):
            jobs.append([NVCC] + FLAGS + ["-DKMI_K=%d" % k, "-c", inst, "-o", o])
    for name in ("kmi_api", "kmi_dist", "kmi_sorted", "kmi_quality"):
        src = os.path.join(CSRC, name + ".cu")
        o = os.path.join(objdir, name + ".o")
        objs.append(o)
        if force or _newer(o, hdrs + [src]):
            jobs.append([NVCC] + FLAGS + ["-c", src, "-o", o])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    if jobs or not os.path.exists(OUT):
        _run([NVCC, "-shared", "-o", OUT] + objs + ["-cudart", "static", "-ldl"])
    if verbose:
        print("built", OUT)
    return OUT


# k-mer lengths whose kernels are built as loadable modules by default (kmods/libkmi_k<K>.so); any K in 2..31, 33..63 can be
# added with --k / --all-k. libkmi.so loads them on demand (kmi_create).
KMODS_DEFAULT = (17, 25, 27, 33, 55)


def build_kmods(ks=KMODS_DEFAULT, force=False, verbose=False):
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".hpp"))]
    inst = os.path.join(CSRC, "kmi_ops_inst.cu")
    outdir = os.path.join(HERE, "kmods")
    os.makedirs(outdir, exist_ok=True)
    jobs, outs = [], []
    for k in ks:
        if k in KS:
            continue
        if not (2 <= k <= 63) or k == 32:
            raise ValueError("k=%d needs >= 2 pad bits (2..31, 33..63)" % k)
        o = os.path.join(outdir, "libkmi_k%d.so" % k)
        outs.append(o)
        if force or _newer(o, hdrs + [inst]):
            jobs.append([NVCC] + FLAGS + ["-DKMI_K=%d" % k, "-DKMI_MODULE", "-shared", inst, "-o", o, "-cudart", "static"])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    if verbose:
        print("k-mer length modules:", ", ".join(os.path.basename(o) for o in outs))
    return outs


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    ks = list(KMODS_DEFAULT)
    if "--all-k" in sys.argv:
        ks = [k for k in range(2, 64) if k != 32]
    if "--k" in sys.argv:
        ks = [int(x) for x in sys.argv[sys.argv.index("--k") + 1:] if x.isdigit()]
    build_kmods(ks, force="--force" in sys.argv, verbose=True)
