"""Build the CUDA C-ABI library in-tree: mantrap_b200/libmantrap_b200.so (sm_100a only)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "libmantrap_b200.so")
SOURCES = [os.path.join(HERE, "csrc", name) for name in ("mtb_kernels.cu", "mtb_jacobian.cu", "mtb_ego.cu", "mtb_scene.cu", "mtb_probe.cu")]
DEPS = SOURCES + [os.path.join(HERE, "csrc", "mtb_math.cuh"), os.path.join(HERE, "csrc", "mtb_common.cuh"), os.path.join(ROOT, "include", "mantrap_b200.h")]
# --use_fast_math: MUFU rsqrt / rcp / ex2 instead of the IEEE division / sqrt / expf sequences (each ~10 instructions
# in the pair term). Parity with the reference is measured WITH this flag (tests -m gpu); NaN / inf semantics that
# define parity (x != x tests, inf - inf) are unaffected by it.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
              "--use_fast_math",
              "-Xcompiler", "-fPIC", "-shared", "-cudart", "static"]


def nvcc() -> str:
    path = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(path):
        raise RuntimeError("nvcc not found: the mantrap_b200 CUDA library cannot be built")
    return path


def build(force: bool = False, verbose: bool = False) -> str:
    extra = os.environ.get("MTB_NVCC_EXTRA", "").split()      # development knob (e.g. --use_fast_math experiments)
    if not force and not extra and os.path.exists(LIB_PATH) and \
            all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in DEPS):
        return LIB_PATH
    cmd = [nvcc(), *NVCC_FLAGS, *extra, "-I", os.path.join(ROOT, "include"), "-I", os.path.join(HERE, "csrc"), "-o", LIB_PATH, *SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        print(proc.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force=True, verbose=True))
