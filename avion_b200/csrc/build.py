"""Compile the C-ABI library in-tree: avion_b200/csrc/libavion_b200.so (nvcc, sm_100a only)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CU = ["api.cu", "gemm.cu", "attention_fwd.cu", "attention_bwd.cu", "transform.cu", "layernorm.cu", "elementwise.cu",
      "head.cu", "infonce.cu", "optim.cu", "videomae.cu", "ranking.cu"]
CPP = ["rrc.cpp"]
LIB = os.path.join(HERE, "libavion_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
         "--use_fast_math" if False else "-DAVN_NO_FAST_MATH"]


def _stale(out, deps):
    return not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps)


def build(force=False, verbose=False, tag=None, extra_flags=()):
    """tag + extra_flags: an experimental variant (kernel A/B runs, tools/ only) built with additional -D flags into
    <repo>/variants/<tag>/libavion_b200.so, objects under build_<tag>/; the product library is the untagged build."""
    global LIB
    objdir = os.path.join(HERE, "build" if not tag else "build_" + tag)
    os.makedirs(objdir, exist_ok=True)
    lib = LIB
    if tag:
        vdir = os.path.join(HERE, "..", "..", "variants", tag)
        os.makedirs(vdir, exist_ok=True)
        lib = os.path.join(vdir, "libavion_b200.so")
    common = [os.path.join(HERE, "common.cuh"), os.path.join(HERE, "..", "..", "include", "avion_b200.h"), __file__]
    jobs = []
    for src in CU + CPP:
        obj = os.path.join(objdir, src + ".o")
        path = os.path.join(HERE, src)
        if force or _stale(obj, [path] + common):
            cmd = [NVCC] + FLAGS + list(extra_flags) + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
            jobs.append(cmd)
    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(run, jobs))
    objs = [os.path.join(objdir, s + ".o") for s in CU + CPP]
    if force or jobs or _stale(lib, objs):
        cmd = [NVCC, "-shared", "-o", lib] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed: %s\n%s" % (r.stdout, r.stderr))
    return lib


if __name__ == "__main__":
    # python build.py [--force] [-v] [--tag NAME -DFLAG ...]
    tag = sys.argv[sys.argv.index("--tag") + 1] if "--tag" in sys.argv else None
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, tag=tag, extra_flags=[a for a in sys.argv[1:] if a.startswith("-D")]))
