"""In-tree build of libboxmot_b200.so (nvcc, sm_100a only).  The .so is git-ignored but
travels to the GPU box with the gpurun snapshot.  Each translation unit is compiled to an
object of its own (in parallel, rebuilt only when it or a header is newer) and linked."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libboxmot_b200.so")
HEADER = os.path.join(HERE, "..", "include", "boxmot_b200.h")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    # fp64 products and sums must round separately, as numpy's do (see csrc/common.cuh)
    "-fmad=false",
    "-Xcompiler", "-fPIC",
]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_mtime():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")] + [HEADER]
    return max(os.path.getmtime(h) for h in hs)


def _obj(src):
    return os.path.join(OBJ, src[:-3] + ".o")


def _stale_objs():
    ht = _headers_mtime()
    out = []
    for s in sources():
        o = _obj(s)
        if not os.path.exists(o) or os.path.getmtime(o) < max(ht, os.path.getmtime(os.path.join(CSRC, s))):
            out.append(s)
    return out


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    todo = sources() if force else _stale_objs()
    objs = [_obj(s) for s in sources()]
    if not todo and os.path.exists(LIB) and all(os.path.getmtime(o) <= os.path.getmtime(LIB) for o in objs):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

    def compile_one(src):
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", os.path.join(CSRC, src), "-o", _obj(src)]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        return src, r.returncode, r.stdout

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(todo)))) as ex:
        results = list(ex.map(compile_one, todo))
    for src, rc, out in results:
        if out.strip() and (verbose or rc != 0 or "warning" in out):
            sys.stderr.write("== %s\n%s\n" % (src, out))
    bad = [src for src, rc, _ in results if rc != 0]
    if bad:
        raise RuntimeError("nvcc failed for: " + ", ".join(bad))
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + objs + ["-o", LIB])
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
