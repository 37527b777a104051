"""TEST INFRASTRUCTURE -- builds the *unmodified* reference rasterizer for sm_100.

Compiles /root/reference/submodules/diff-surfel-rasterization (the seven sources its
setup.py:24-31 lists) from where they lie, with nvcc directly (the reference's own build
system is not run).  gcc 13 needs <cstdint> in cuda_rasterizer/rasterizer_impl.h; rather
than patching the source it is force-included (`-include cstdint`).  Outputs go only to
oracle/_ref/ (git-ignored, travels to the GPU box):

    oracle/_ref/diff_surfel_rasterization/_C.so        compiled extension (sm_100)
    oracle/_ref/diff_surfel_rasterization/__init__.py  installed copy of the reference's python
                                                       wrapper (what `pip install --target` would
                                                       place there; never committed)

Only tests/, __graft_entry__ and bench.py (reference arm / parity check) load the result.
"""
import os
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/submodules/diff-surfel-rasterization"
OUT = os.path.join(HERE, "_ref", "diff_surfel_rasterization")
OBJ = os.path.join(HERE, "_ref", "obj")
SOURCES = [
    "cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
    "cuda_rasterizer/backward.cu", "cuda_rasterizer/adam.cu",
    "rasterize_points.cu", "conv.cu", "ext.cpp",
]


def main() -> int:
    if not os.path.isdir(SRC):
        print("reference sources absent; keeping any prebuilt oracle/_ref", file=sys.stderr)
        return 0
    so = os.path.join(OUT, "_C.so")
    newest = max(os.path.getmtime(os.path.join(SRC, s)) for s in SOURCES)
    if os.path.exists(so) and os.path.getmtime(so) > newest and "--force" not in sys.argv:
        return 0
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    inc = ["-I" + p for p in ce.include_paths(device_type="cuda")]
    inc += ["-I" + sysconfig.get_paths()["include"], "-I" + os.path.join(SRC, "third_party/glm")]
    common = ["-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
              "-D_GLIBCXX_USE_CXX11_ABI=1", "-std=c++17", "-O3"]
    # same code-generation defaults as the reference's setup.py (no fast-math, fmad on), arch sm_100
    nvcc = ["nvcc", "-gencode", "arch=compute_100,code=sm_100", "-w", "--expt-relaxed-constexpr",
            "-include", "cstdint", "-Xcompiler", "-fPIC", "-lineinfo"]

    def compile_one(s):
        o = os.path.join(OBJ, s.replace("/", "_") + ".o")
        if s.endswith(".cu"):
            cmd = nvcc + common + inc + ["-c", os.path.join(SRC, s), "-o", o]
        else:
            cmd = ["g++", "-fPIC", "-include", "cstdint"] + common + inc + ["-c", os.path.join(SRC, s), "-o", o]
        subprocess.run(cmd, check=True)
        return o

    with ThreadPoolExecutor(max_workers=7) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    libdirs = ce.library_paths(device_type="cuda")
    link = ["g++", "-shared", "-o", so] + objs + ["-L" + p for p in libdirs] + \
           ["-Wl,-rpath," + p for p in libdirs] + \
           ["-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
    subprocess.run(link, check=True)
    shutil.copyfile(os.path.join(SRC, "diff_surfel_rasterization/__init__.py"), os.path.join(OUT, "__init__.py"))
    print("built", so)
    return 0


if __name__ == "__main__":
    sys.exit(main())
