import sys, os, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import induction_b200
from induction_b200._lib import LIB, ptr, current_stream
T, u4 = 76800, 512
A = torch.randn(T, 128, device="cuda").bfloat16()
X = torch.randn(T, 64, device="cuda").bfloat16()
B = torch.randn(T, u4, device="cuda").bfloat16()
C = torch.zeros(188, u4, device="cuda")
def run(M, Aop, lda, sk):
    LIB.call("indnet_tc_gemm", current_stream(), 1, 1, M, u4, T, ptr(Aop), lda, ptr(B), u4, ptr(C), None, u4, None, 0, 0, 1, 0, 0, 0, sk)
for M, Aop, lda, name in ((128, A, 128, "dWh"), (60, X, 64, "dWx")):
    for sk in (0, 148, 74, 37, 18, 9):
        for _ in range(3): run(M, Aop, lda, sk)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run(M, Aop, lda, sk)
        e1.record(); torch.cuda.synchronize()
        print(name, "splitk", sk, "%.1f us" % (e0.elapsed_time(e1) / 20 * 1e3))

# stacked [x | h_prev]^T dz kernel (indnet_tc_gemm_lstm_dw)
out = torch.zeros(60 + 128, u4, device="cuda")
def run_dual(sk):
    LIB.call("indnet_tc_gemm_lstm_dw", current_stream(), ptr(X), ptr(A), ptr(B), ptr(out), None, T, 60, 128, sk, 0)
for sk in (0, 18, 37, 74, 111, 148):
    for _ in range(3): run_dual(sk)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): run_dual(sk)
    e1.record(); torch.cuda.synchronize()
    print("stacked dW splitk", sk, "%.1f us" % (e0.elapsed_time(e1) / 20 * 1e3))
