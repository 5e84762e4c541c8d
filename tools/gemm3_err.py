"""Error of the split-bf16 GEMM (indnet_tc_gemm3) and of cuBLAS SGEMM against float64, by K."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_tc_gemm_gpu import _split3
from induction_b200._lib import LIB, current_stream, ptr
torch.backends.cuda.matmul.allow_tf32 = False
M, N = 192, 512
for K in (64, 256, 1024, 4096, 16384, 76800):
    for sk in (1, 0):
        g = torch.Generator(device="cuda").manual_seed(K)
        A = torch.randn(K, M, generator=g, device="cuda")
        B = torch.randn(K, N, generator=g, device="cuda")
        Ah, Am, Al = _split3(A, M)
        Bh, Bm, Bl = _split3(B, N)
        C = torch.zeros(M, N, device="cuda")
        LIB.call("indnet_tc_gemm3", current_stream(), 1, 1, M, N, K, ptr(Ah), ptr(Am), ptr(Al), M,
                 ptr(Bh), ptr(Bm), ptr(Bl), N, ptr(C), N, None, 0, 0, 1, 0, 0, M * N, sk)
        torch.cuda.synchronize()
        want = A.double().t() @ B.double()
        ref = A.t() @ B
        d = (C.double() - want)
        d2 = (ref.double() - want)
        sc = float(want.abs().max())
        print("K %6d splitk %d: tc3 max %.2e mean-signed %.2e (sign-corr %.2e) | sgemm max %.2e" % (
            K, sk, float(d.abs().max()) / sc, float(d.mean()) / sc,
            float((d * want.sign()).mean()) / sc, float(d2.abs().max()) / sc), flush=True)
