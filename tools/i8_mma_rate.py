#!/usr/bin/env python
"""diagnostic: rate of tcgen05.mma kind::i8 (M = 128, K = 32) as a function of the instruction's N (tb200_i8_mma_rate):
decides how the stacked-slice kernels cut their B operand into instructions.  Prints a markdown table."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from turbo_b200 import _lib  # noqa: E402

lib = _lib.load()
torch.cuda.set_device(0)
print("| N | TOP/s, zero operands | cycles per MMA at 1.90 GHz (zero operands) | TOP/s, pseudo-random digits |")
print("|---|---|---|---|")
sms = torch.cuda.get_device_properties(0).multi_processor_count
for n in (16, 32, 64, 96, 128, 160, 192, 224, 256):
    out = []
    for rnd in (0, 1):
        v = C.c_double()
        _lib.check(lib.tb200_i8_mma_rate(20.0, rnd, n, C.byref(v)))
        out.append(v.value)
    cyc = 2.0 * 128 * n * 32 * sms * 1.90e9 / (out[0] * 1e12)
    print("| %d | %.0f | %.1f | %.0f |" % (n, out[0], cyc, out[1]))
