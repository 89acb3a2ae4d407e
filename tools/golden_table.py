"""Per-fixture parity table: for every golden scene the raw costs the reference recorded, the size of the
reference's gradients and the relative error this implementation achieves against them.

    python tools/golden_table.py            # CUDA library on cuda:0 (GPU box)
    python tools/golden_table.py --emul     # host emulation of the kernel sources (build container)
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_io as gio  # noqa: E402
import kernel_cases as kc  # noqa: E402

if __name__ == "__main__":
    if "--emul" in sys.argv:
        sys.path.insert(0, os.path.join(ROOT, "tests", "host_emul"))
        import build_emul
        from mantrap_b200 import _cabi
        lib, dev = _cabi.bind(ctypes.CDLL(build_emul.build())), "cpu"
    else:
        from mantrap_b200 import engine
        lib, dev = engine.Engine("cuda:0").lib, "cuda:0"
    print("| fixture | min sigma | module | raw cost | max abs(grad_ref) | obj rel. err | grad rel. err |")
    print("|---|---|---|---|---|---|---|")
    for name in gio.case_names():
        g = gio.load(name)
        worst = kc.check_modules(lib, dev, name)
        smin = f"{float(min(g['sigma'].min(), g['sigma_wo'].min())):.3f}" if str(g["env"]) == "sf" else "-"
        for key in ("acc", "vel", "pos", "prob"):
            if key not in worst:
                continue
            raw = float(g[f"raw_{key}"])
            gmax = float(np.abs(g[f"grad_{key}"]).max())
            e_obj, e_grad = worst[key]
            print(f"| {name} | {smin} | {key} | {raw:.4g} | {gmax:.3g} | {e_obj:.1e} | "
                  f"{(f'{e_grad:.1e}' if gmax > 0 else '0 = 0')} |")
