"""Dev: phase clocks of the Gram ndiv kernels (needs the -DNDP_TC_PROF build: tools/build_prof.sh, run with
NDP_LIB_PATH=ndivplanning_b200/libndivplan_prof.so).  Row blocks emulate what a rank of a P-way sharded call runs."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import diversity as div, synthetic  # noqa: E402
from ndivplanning_b200._lib import lib  # noqa: E402

NAMES = ["ctas", "prologue", "first operands", "main loop", "results+row reduce", "grid tail", "max total"]


def main():
    L = lib()
    L.ndp_debug_gram_prof.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
    out = (C.c_ulonglong * 16)()
    x, z = synthetic.make_ndiv_inputs(1, 4096, 32, 16)
    x, z = x.cuda(), z.cuda()
    sx, sz = div.ndiv_rowsums(x, z)
    for rows in (4096, 2048, 512):
        for _ in range(3):
            div.ndiv_rowsums(x, z, 0, rows, sx.clone(), sz.clone())
            div.ndiv_loss_grad(x, z, sx, sz, True, 0, rows)
        L.ndp_debug_gram_prof(out, 1)
        reps = 10
        for _ in range(reps):
            div.ndiv_rowsums(x, z, 0, rows, sx.clone(), sz.clone())
            div.ndiv_loss_grad(x, z, sx, sz, True, 0, rows)
        L.ndp_debug_gram_prof(out, 1)
        for ps in (0, 1):
            v = list(out)[8 * ps:8 * ps + 8]
            n = max(v[0], 1)
            print(f"rows {rows} pass {ps + 1}: {v[0] // reps} CTAs/launch; mean cycles per CTA: " +
                  ", ".join(f"{NAMES[i]} {v[i] / n:.0f}" for i in range(1, 6)) + f"; max total {v[6]}")


if __name__ == "__main__":
    main()
