import torch, sys, os
from cmf_b200 import kernels as K
N, H, W = 3250, 28, 28
x = torch.randn(N, 64, H, W, device="cuda")
w = torch.randn(64, 64, 3, 3, device="cuda") / 24
ws = K.conv3x3_c64_prep(w, False)
for _ in range(3):
    K.conv3x3_c64(x, ws, None)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    K.conv3x3_c64(x, ws, None)
e1.record()
torch.cuda.synchronize()
print("dbg", os.environ.get("CMF_CONV_DBG", "0"), "ms", e0.elapsed_time(e1) / 10, flush=True)
import ctypes, numpy as np
from cmf_b200 import _lib
lib = _lib.load()
buf = (ctypes.c_longlong * 128)()
ctypes.CDLL(_lib.LIB_PATH).cmf_debug_conv_prof(buf)
a = np.array(buf[:], dtype=np.int64)
t0 = a[0]
print("setup", a[1] - t0)
print("worker pre-wait ", (a[64:82] - t0).tolist())
print("worker post-wait", (a[96:114] - t0).tolist())
print("worker arrive   ", (a[2:20] - t0).tolist())
print("mma issued      ", (a[32:50] - t0).tolist())
print("epilogue end", a[30] - t0)
print("mma stage 2 detail", (a[50:58] - t0).tolist())
