"""Build the CUDA C-ABI library in-tree: mantrap_b200/libmantrap_b200.so (sm_100a only)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "libmantrap_b200.so")
# test-only variant with the launch-geometry development knobs (-DMTB_DEV_KNOBS): the production library never reads the
# environment; two placement tests load this one through Engine.from_library
DEV_LIB_PATH = os.path.join(HERE, "libmantrap_b200_dev.so")
SOURCES = [os.path.join(HERE, "csrc", name) for name in ("mtb_kernels.cu", "mtb_jacobian.cu", "mtb_ego.cu", "mtb_scene.cu", "mtb_probe.cu")]
DEPS = sorted(os.path.join(HERE, "csrc", name) for name in os.listdir(os.path.join(HERE, "csrc"))) + [os.path.join(ROOT, "include", "mantrap_b200.h")]
# --use_fast_math: MUFU rsqrt / rcp / ex2 instead of the IEEE division / sqrt / expf sequences (each ~10 instructions
# in the pair term). Parity with the reference is measured WITH this flag (tests -m gpu); NaN / inf semantics that
# define parity (x != x tests, inf - inf) are unaffected by it.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
              "--use_fast_math",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fno-gnu-unique", "-shared", "-cudart", "static"]
# -fno-gnu-unique: function-local statics of templates would otherwise get STB_GNU_UNIQUE binding, which the dynamic loader
# resolves process-wide even for RTLD_LOCAL libraries -- a process that also loads the test-only emulation build of the same
# sources (tests/host_emul) would then share launch tables between the two libraries.


def nvcc() -> str:
    path = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(path):
        raise RuntimeError("nvcc not found: the mantrap_b200 CUDA library cannot be built")
    return path


def build(force: bool = False, verbose: bool = False, dev_knobs: bool = False) -> str:
    out = DEV_LIB_PATH if dev_knobs else LIB_PATH
    extra = ["-DMTB_DEV_KNOBS"] if dev_knobs else []
    if not force and os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in DEPS):
        return out
    cmd = [nvcc(), *NVCC_FLAGS, *extra, "-I", os.path.join(ROOT, "include"), "-I", os.path.join(HERE, "csrc"), "-o", out, *SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        print(proc.stderr)
    return out


if __name__ == "__main__":
    print(build(force=True, verbose=True))
