"""Per-step timeline of CTA (0,0) of tc_lstm_fwd_kernel.  Needs the timeline build:
    INDNET_NVCC_DEFINES=-DINDNET_LSTM_TIMELINE python induction-network-on-fewrel_b200/csrc/build.py --force
(rebuild without the define afterwards)."""
import sys, os, ctypes, runpy
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import numpy as np, torch
import induction_b200
from induction_b200._lib import LIB
sys.argv = ["bench.py", "--steps", "2", "--warmup", "3", "--no-cpu-baseline", "--no-graph"]
try:
    runpy.run_path(os.path.join(root, "bench.py"), run_name="__main__")
except SystemExit:
    pass
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 512)()
LIB.load()
LIB._dll.indnet_debug_lstm.argtypes = [ctypes.c_void_p]
rc = LIB._dll.indnet_debug_lstm(buf)
a = np.array(list(buf), dtype=np.int64).reshape(64, 8)
print("rc", rc)
print("forward, gate thread 0, per step and half: wait for MMA | tmem ld | math + st.shared | fence + arrive   (A || B) || period")
for s in range(1, 38):
    r, n, pv = a[s], a[s + 1], a[s - 1]
    print(s, r[0] - pv[7], '|', r[1] - r[0], '|', r[2] - r[1], '|', r[3] - r[2], '||', r[4] - r[3], '|', r[5] - r[4], '|', r[6] - r[5], '|', r[7] - r[6],
          '|| period', n[0] - r[0])
rc = LIB._dll.indnet_debug_lstm_bwd(buf)
b = np.array(list(buf), dtype=np.int64).reshape(64, 8)
print("BPTT rc", rc)
print("gate thread 0, per step and half: wait inputs | lds + dropout + wait MMA + tmem ld | math + st.shared | dx store + fence + arrive  (A || B) || period")
for s in range(2, 38):
    r, n, pv = b[s], b[s + 1], b[s - 1]
    print(s, r[0] - pv[7], '|', r[1] - r[0], '|', r[2] - r[1], '|', r[3] - r[2], '||', r[4] - r[3], '|', r[5] - r[4], '|', r[6] - r[5], '|', r[7] - r[6],
          '|| period', n[0] - r[0])
