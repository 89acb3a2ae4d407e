"""Build the host-emulation library of the kernel source (test infrastructure, see cuda_shim.h)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "libmtb_emul.so")
SOURCES = [os.path.join(ROOT, "mantrap_b200", "csrc", f) for f in ("mtb_kernels.cu", "mtb_jacobian.cu", "mtb_ego.cu", "mtb_scene.cu")]
CSRC = os.path.join(ROOT, "mantrap_b200", "csrc")
DEPS = sorted(os.path.join(CSRC, name) for name in os.listdir(CSRC)) + [os.path.join(HERE, "cuda_shim.h"), os.path.join(ROOT, "include", "mantrap_b200.h")]


def build(force: bool = False) -> str:
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in DEPS):
        return OUT
    cmd = ["g++", "-std=c++20", "-O2", "-g", "-DMTB_HOST_EMUL", "-DMTB_DEV_KNOBS", "-fPIC", "-shared", "-pthread", "-ffp-contract=off", "-fno-gnu-unique",
           "-I", HERE, "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "mantrap_b200", "csrc"), "-o", OUT]
    for src in SOURCES:
        cmd += ["-x", "c++", src]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
