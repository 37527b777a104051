"""Builds libsurfel_b200.so (hand-written sm_100a CUDA behind the C ABI of include/surfel_b200.h).

In-tree build: objects under vings-mono_b200/csrc/build/, the shared library next to the Python
shim (vings-mono_b200/diff_surfel_rasterization/libsurfel_b200.so) so that it travels to the GPU box.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
# SB200_VARIANT=<name> + SB200_EXTRA_FLAGS="-D..." build a side-by-side debug library (loaded with SB200_LIB)
VARIANT = os.environ.get("SB200_VARIANT", "")
OBJ = OBJ + ("_" + VARIANT if VARIANT else "")
OUT = os.path.join(HERE, "diff_surfel_rasterization", "libsurfel_b200%s.so" % ("_" + VARIANT if VARIANT else ""))
SOURCES = ["api.cu", "preprocess.cu", "binning.cu", "radix_sort.cu", "blend_forward.cu", "blend_backward.cu",
           "preprocess_backward.cu", "map_loss.cu", "surfel_update.cu", "dp_exchange.cu", "churn.cu", "sh_color.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v"] + os.environ.get("SB200_EXTRA_FLAGS", "").split()


def csrc_hash():
    """sha1 over the kernel sources and the ABI header with comments and whitespace stripped: stamps
    profiles/traffic.json (tools/traffic_from_ncu.py) so that bench.py can tell when an ncu capture predates the kernels
    it describes -- a comment-only edit does not invalidate a capture, any change of code does."""
    import hashlib
    import re
    strip = re.compile(r"/\*.*?\*/|//[^\n]*", re.S)
    h = hashlib.sha1()
    files = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh")))
    paths = [os.path.join(CSRC, f) for f in files] + [os.path.join(HERE, "..", "include", "surfel_b200.h")]
    for path in paths:
        h.update(os.path.basename(path).encode())
        h.update("".join(strip.sub("", open(path).read()).split()).encode())
    return h.hexdigest()


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(HERE, "..", "include", "surfel_b200.h"))

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        if force or _stale(o, [s] + headers):
            r = subprocess.run(["nvcc"] + NVCC_FLAGS + ["-c", s, "-o", o], capture_output=True, text=True)
            with open(o + ".log", "w") as f:
                f.write(r.stdout + r.stderr)
            if r.returncode != 0:
                raise RuntimeError("nvcc failed on %s:\n%s" % (src, r.stdout + r.stderr))
            if verbose:
                print(r.stderr)
        return o

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(OUT, objs):
        # (the arch is repeated at the link step: without it nvcc compiles its link stub for its default sm_52)
        subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", OUT] + objs + ["-lcudart"],
                       check=True)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
