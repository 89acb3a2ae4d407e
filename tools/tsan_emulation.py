"""Race check of the kernel sources on the CPU: builds the host emulation (every CUDA thread = one OS thread, barriers =
std::barrier, shuffles through per-warp buffers) with -fsanitize=thread and drives every kernel on small shapes. A missing
barrier / a shared-memory hazard in the CUDA source shows up as a ThreadSanitizer data race here (compute-sanitizer's
racecheck is not available on the GPU pool). TEST INFRASTRUCTURE, like the emulation itself.

    LD_PRELOAD=$(gcc -print-file-name=libtsan.so) TSAN_OPTIONS="report_signal_unsafe=0 exitcode=0" python tools/tsan_emulation.py
"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
EMUL = os.path.join(ROOT, "tests", "host_emul")
OUT = os.path.join(EMUL, "libmtb_emul_tsan.so")


def build(selftest=False):
    srcs = [os.path.join(ROOT, "mantrap_b200", "csrc", f) for f in ("mtb_kernels.cu", "mtb_jacobian.cu", "mtb_ego.cu", "mtb_scene.cu")]
    cmd = ["g++", "-std=c++20", "-O1", "-g", "-fsanitize=thread", "-DMTB_HOST_EMUL", "-fPIC", "-shared", "-pthread", "-ffp-contract=off",
           "-I", EMUL, "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "mantrap_b200", "csrc"), "-o", OUT]
    if selftest:
        cmd.insert(1, "-DMTB_EMUL_BREAK_GROUP_BARRIER")
    for s in srcs:
        cmd += ["-x", "c++", s]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    if "libtsan" not in os.environ.get("LD_PRELOAD", ""):
        sys.exit("run with LD_PRELOAD=$(gcc -print-file-name=libtsan.so)")
    import torch
    from mantrap_b200 import _cabi, engine, mpc, scenarios
    torch.set_num_threads(1)
    selftest = "--selftest" in sys.argv      # group barrier removed: the run must produce ThreadSanitizer reports
    eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build(selftest))), "cpu")
    shapes = (("sf", 6, 3, 5), ("sf", 12, 3, 4), ("pf", 5, 5, 4), ("sf", 2, 2, 6), ("kalman", 3, 1, 4))
    for env_kind, n, p, t_hor in (shapes[1:2] if selftest else shapes):
        s, c = 2, 3
        scene = scenarios.synthetic_scenes(s, n, p, env_kind=env_kind, regular_sigma=True)
        controls = scenarios.candidate_controls(s, c, t_hor)
        mus_wo = eng.rollout_wo_ego(scene, t_hor)["mus"].contiguous()
        spec = engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0, "prob": 1.0}, mus_wo=mus_wo, control_limits=True, speed_limits=True,
                                 min_distance=True, ellipsoid=True)
        out = eng.evaluate(scene, controls, spec, want=("obj", "total", "grad", "cons", "violation", "mus", "sigmas"))
        eng.argmin(out["total"], out["violation"], controls)
        eng.constraint_jacobian(scene, controls, spec)
        eng.attention("closest", scene.ego0, scene.ado0)
        mpc.BatchedClosedLoop(eng, scene, t_planning=t_hor, num_candidates=c, attention="euclidean").solve(1)
        print("ok", env_kind, n, p, t_hor, float(out["total"].mean()), flush=True)
