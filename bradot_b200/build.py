"""Build libbradot_physics_cuda.so in-tree with nvcc for sm_100a (no JIT cache, the .so travels)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libbradot_physics_cuda.so")
SOURCES = ["bpc_api.cu", "bpc_bodies.cu", "bpc_step.cu", "k_sort.cu", "k_integrate.cu", "k_broadphase.cu", "k_graph.cu",
           "k_narrowphase.cu", "k_solver.cu", "k_queries.cu", "k_query_shapes.cu", "k_areas.cu"]
# -fmad=false: the reference is built with -ffp-contract=off; structural comparisons must agree bit for bit
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-O2"]


def _stale(obj, src):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    deps = [src] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    deps.append(os.path.join(HERE, "..", "include", "bradot_physics_cuda.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose=False, force=False, lib=LIB, defines=()):
    """`lib` / `defines`: A/B builds of the same ABI next to the default one (BPC_LIBRARY selects one at run time)."""
    nvcc = os.environ.get("NVCC", "nvcc")
    objs = []
    procs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(CSRC, s.replace(".cu", ".o" if lib == LIB else "." + os.path.basename(lib) + ".o"))
        objs.append(obj)
        if force or _stale(obj, src):
            cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write("nvcc failed for %s:\n%s\n" % (s, out))
        elif verbose:
            print(out)
    if failed:
        raise RuntimeError("CUDA build failed")
    if procs or not os.path.exists(lib):
        subprocess.run([nvcc, "-shared", "-o", lib] + objs + ["-lcudart"], check=True)
    return lib


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
