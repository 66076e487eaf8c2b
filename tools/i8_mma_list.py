#!/usr/bin/env python
"""diagnostic: SM cycles per K = 32 step of compile-time MMA lists issued back to back on resident operands
(tb200_i8_mma_list, csrc/ozaki_i8.cuh MmaList<V>).  The variance kernel at 21 slice products spends ~865 cycles per K = 32
step where its instruction list should cost 721 by the per-N rate table (profiles/r2_i8_mma_rate.md) and 672 by its
multiply-accumulates; this asks the tensor pipe which property of the list costs the difference:
    python tools/i8_mma_list.py > gpurun_out/i8_mma_list.md"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from turbo_b200 import _lib  # noqa: E402

NAMES = [
    "variance kernel, 21 products: plane i x stacked planes 1..7-i, balanced chunks (192+192, 160+160, 256, 192, 128, 64)",
    "the same instructions with ONE A plane",
    "... one A plane and B rows from 0",
    "the kernel's operands and N sequence, accumulators alternating between the aligned halves of TMEM",
    "uniform: 7 x N=192, planes in turn, alternating halves",
    "the kernel's list, every accumulator range starting at column 0",
    "the kernel's list, short instructions first",
    "chunks of 256 + remainder (256+128, 256+64, 256, 192, 128, 64)",
    "variance kernel, 28 products (224+224, 192+192, 160+160, 256, 192, 128, 64)",
    "21 products, short and long instructions interleaved",
    "uniform: 6 x N=224, alternating halves",
    "5 x N=256 + 1 x N=64, alternating halves",
    "the kernel's list, A kept in the collector for the second chunk of a range (collector::a::fill / lastuse)",
    "... and short / long instructions interleaved",
]


def run(lib, v, planes):
    cyc, n, cols = C.c_double(), C.c_int(), C.c_int()
    _lib.check(lib.tb200_i8_mma_list(v, 4000, planes, C.byref(cyc), C.byref(n), C.byref(cols)))
    return cyc.value / 2, n.value, cols.value


def main():
    lib = _lib.load()
    torch.cuda.set_device(0)
    print("cycles per K = 32 step; resident = operands stay in shared memory; pipeline p = two-stage pipeline whose idle "
          "stage is refilled with p digit planes of both operands per k-block (6 = the variance kernel at 21 products: "
          "36 KB per step)")
    print()
    print("| # | list | MMAs per step | ideal (columns / 2) | resident | pipeline 0 | pipeline 2 | pipeline 4 | pipeline 6 | pipeline 7 |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for v, name in enumerate(NAMES):
        res, n, cols = run(lib, v, -1)
        row = [run(lib, v, p)[0] for p in (0, 2, 4, 6, 7)] if v in (0, 4, 8, 9, 10, 12, 13) else [float('nan')] * 5
        print("| %d | %s | %d | %.0f | %.0f | %s |" % (v, name, n, cols / 2, res, " | ".join("%.0f" % r for r in row)))


if __name__ == "__main__":
    main()
