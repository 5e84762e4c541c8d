"""GPU: ``indnet_lstm_bwd_static`` - the BPTT entry point for callers that cannot allocate device
memory (custom-op / plugin callbacks handed only their input and output buffers, loops that must not
hit cudaMalloc's device-wide synchronisation).  Its scratch is a fixed 512 KiB ``__device__`` array
inside the library instead of a caller-provided buffer, so it is limited to u <= 128.

Checked here, through the C-ABI: on two calls in a row (the second with other weights and a smaller
hidden size, so stale contents of the shared array from the first call are in play) its ``dz`` output
matches ``indnet_lstm_bwd`` within the gradient tolerance the other GPU tests use; sizes beyond the
built-in maximum are refused with a clear message; the call leaves the amount of free device memory
unchanged.  This file sorts last on purpose: a run that stops at the first failure has then already
run every other test.  It does not depend on them and also passes alone in a fresh process.
"""
import pytest
import torch

from _cases import rel_err

pytestmark = pytest.mark.gpu

GRAD_TOL = 2e-4   # same bound as tests/test_pipeline_gpu.py


def _inputs(X, L, u, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g, device="cuda")          # noqa: E731
    n = lambda *s: torch.randn(*s, generator=g, device="cuda")         # noqa: E731
    gates = r(2, X * L, 4 * u)
    gates[:, :, u:2 * u] = gates[:, :, u:2 * u] * 2 - 1                # j gate is a tanh output
    return dict(dhout=n(X * L, 2 * u) * 0.1, gates=gates.contiguous(), cstate=n(2, X * L, u),
                wh_fw=n(u, 4 * u) * 0.1, wh_bw=n(u, 4 * u) * 0.1)


def _run_both(X, L, u, seed, keep_prob=1.0):
    from induction_b200._lib import LIB, current_stream, ptr
    t = _inputs(X, L, u, seed)
    scratch = torch.empty(2, 4 * u, u, device="cuda")
    dz_ref = torch.full((2, X * L, 4 * u), float("nan"), device="cuda")
    dz_new = torch.full((2, X * L, 4 * u), float("nan"), device="cuda")
    LIB.call("indnet_lstm_bwd", current_stream(), ptr(t["dhout"]), ptr(t["gates"]), ptr(t["cstate"]),
             ptr(t["wh_fw"]), ptr(t["wh_bw"]), ptr(scratch), ptr(dz_ref), X, L, u, keep_prob, 77, None)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    LIB.call("indnet_lstm_bwd_static", current_stream(), ptr(t["dhout"]), ptr(t["gates"]), ptr(t["cstate"]),
             ptr(t["wh_fw"]), ptr(t["wh_bw"]), ptr(dz_new), X, L, u, keep_prob, 77, None)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return dz_ref, dz_new, free_before, free_after


def test_static_scratch_entry_matches_on_two_calls_in_a_row():
    # call 1: the largest supported hidden size, fills the whole built-in array
    dz_ref, dz_new, _, _ = _run_both(X=37, L=9, u=128, seed=1)
    assert torch.isfinite(dz_new).all()
    assert rel_err(dz_new, dz_ref) < GRAD_TOL
    # call 2, straight after: other weights, smaller u, dropout on - the array still holds call 1's data
    dz_ref, dz_new, free_before, free_after = _run_both(X=18, L=5, u=20, seed=2, keep_prob=0.9)
    assert torch.isfinite(dz_new).all()
    assert rel_err(dz_new, dz_ref) < GRAD_TOL
    assert free_after == free_before, "the no-allocation entry point changed free device memory"


def test_static_scratch_entry_rejects_sizes_beyond_its_maximum():
    from induction_b200._lib import LIB, IndnetError, current_stream, ptr
    X, L, u = 4, 3, 129
    t = _inputs(X, L, u, seed=3)
    dz = torch.empty(2, X * L, 4 * u, device="cuda")
    with pytest.raises(IndnetError, match=r"exceeds the built-in scratch \(max u=128\)"):
        LIB.call("indnet_lstm_bwd_static", current_stream(), ptr(t["dhout"]), ptr(t["gates"]), ptr(t["cstate"]),
                 ptr(t["wh_fw"]), ptr(t["wh_bw"]), ptr(dz), X, L, u, 1.0, 0, None)
