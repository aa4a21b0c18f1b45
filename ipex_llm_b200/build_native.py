"""Build libb200lowbit.so in-tree with nvcc for sm_100a.

    python -m ipex_llm_b200.build_native [--force] [--verbose]

The .so lands in ipex_llm_b200/libs/ (git-ignored, travels with gpurun snapshots).  nvcc
cross-compiles without a GPU.  Objects are cached in ipex_llm_b200/csrc/build/ keyed by the
source + header mtimes so that re-runs are cheap.
"""
import concurrent.futures
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIBDIR = os.path.join(HERE, "libs")
LIB = os.path.join(LIBDIR, "libb200lowbit.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    # IEEE division / sqrt and no flush-to-zero: the quantiser must match the CPU statement
    "--prec-div=true", "--prec-sqrt=true", "--ftz=false",
    "-I", INCLUDE, "-I", CSRC,
]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: libb200lowbit.so cannot be built")
    return exe


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _newest_header():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE) if f.endswith(".h")]
    return max(os.path.getmtime(h) for h in hs)


def _compile(src, verbose, extra):
    obj = os.path.join(OBJ, src[:-3] + ".o")
    cmd = [_nvcc()] + NVCC_FLAGS + extra + os.environ.get("LOWBIT_NVCC_EXTRA", "").split() + ["-c", os.path.join(CSRC, src), "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    return obj, r.stderr


def build(force=False, verbose=False, ptxas_info=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    hdr_time = _newest_header()
    todo, objs = [], []
    for s in sources():
        obj = os.path.join(OBJ, s[:-3] + ".o")
        objs.append(obj)
        src_time = max(os.path.getmtime(os.path.join(CSRC, s)), hdr_time)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < src_time:
            todo.append(s)
    extra = ["-Xptxas", "-v"] if ptxas_info else []
    logs = {}
    if todo:
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            for s, (obj, err) in zip(todo, ex.map(lambda s: _compile(s, verbose, extra), todo)):
                logs[s] = err
    if todo or not os.path.exists(LIB) or force:
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                         "-Xcompiler", "-fPIC", "-cudart", "static"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if ptxas_info:
        for s, err in logs.items():
            print(f"==== {s}\n{err}")
    return LIB


if __name__ == "__main__":
    lib = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv,
                ptxas_info="--ptxas-info" in sys.argv)
    print(lib)
