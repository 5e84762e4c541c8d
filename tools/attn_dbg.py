"""Per-tile timeline of CTA 0 of attn_fused_bwd_kernel (build with INDNET_NVCC_DEFINES=-DINDNET_ATTN_TIMELINE)."""
import sys, os, ctypes, runpy
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import numpy as np, torch
import induction_b200
from induction_b200._lib import LIB
sys.argv = ["bench.py", "--config", "2", "--steps", "2", "--warmup", "3", "--no-cpu-baseline", "--no-graph"]
try:
    runpy.run_path(os.path.join(root, "bench.py"), run_name="__main__")
except SystemExit:
    pass
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 128)()
LIB._dll.indnet_debug_attn.argtypes = [ctypes.c_void_p]
print("rc", LIB._dll.indnet_debug_attn(buf))
a = np.array(list(buf), dtype=np.int64).reshape(16, 8)
print("tile: wait prefetch+sync | phase A | prefetch issue + dscore | wait proj TMA | phase B | fence+sync | MMA+wait | (epilogue -> next top)")
for t in range(5):
    r = a[t]
    nxt = a[t + 1][0] if a[t + 1][0] else 0
    print(t, r[1]-r[0], '|', r[2]-r[1], '|', r[3]-r[2], '|', r[4]-r[3], '|', r[5]-r[4], '|', r[6]-r[5], '|', r[7]-r[6], '|', (nxt - r[7]) if nxt else -1)

if hasattr(LIB._dll, "indnet_debug_attn_fwd"):
    LIB._dll.indnet_debug_attn_fwd.argtypes = [ctypes.c_void_p]
    print("fwd rc", LIB._dll.indnet_debug_attn_fwd(buf))
    a = np.array(list(buf), dtype=np.int64).reshape(16, 8)
    print("forward (persistent), CTA 0, epilogue warp 2 lane 0, per tile: wait acc | drain | bar | softmax (4 bars) | pooling | end bar | (-> next tile top)")
    for t in range(5):
        r = a[t]
        nxt = a[t + 1][0] if a[t + 1][0] else 0
        print(t, r[1]-r[0], '|', r[2]-r[1], '|', r[3]-r[2], '|', r[4]-r[3], '|', r[5]-r[4], '|', r[6]-r[5], '|', (nxt - r[6]) if nxt else -1)
