"""Per-phase clock64() breakdown of the backward kernel.  Needs a development build:
   ENARF_NVCC_EXTRA=-DENARF_PHASE_TIMING python -c "from enarf_gan_b200.build import build_library as b; b(force=True)" """
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from enarf_gan_b200 import _abi
dev = torch.device("cuda:0"); torch.cuda.set_device(0)
w = bench.build_workload(list(range(8)), dev)
for i in range(2): bench.step_device(w, i)
torch.cuda.synchronize()
lib = _abi.lib()
buf = (ctypes.c_ulonglong * 16)()
lib.enarf_debug_bwd_phases(buf, 1)
n = 3
for i in range(n): bench.step_device(w, 10 + i)
torch.cuda.synchronize()
lib.enarf_debug_bwd_phases(buf, 1)
names = ["tile load", "tile_points", "composite bwd", "records + X load", "L1", "L2", "head", "G2 pass", "dW2", "G1", "dW1", "dX",
         "scatter", "final flush", "-", "tile-top barrier"]
tot = sum(buf)
for nm, v in zip(names, buf):
    if v: print("%-18s %6.2f%%   %.3f ms/CTA/launch" % (nm, 100.0 * v / tot, v / n / 296 / 1.965e6))
print("total %.3f ms per CTA per launch" % (tot / n / 296 / 1.965e6))
