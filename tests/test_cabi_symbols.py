"""CPU: the C-ABI library loads and exports every symbol include/indnet.h declares (no compute
calls - there is no GPU here), argument errors come back as codes + messages, and the product fails
loudly without a device instead of falling back."""
import ctypes
import os

import pytest
import torch

import induction_b200  # noqa: F401
from induction_b200 import _lib


def test_header_parses_and_library_exports_every_symbol():
    protos = _lib.parse_header()
    assert len(protos) >= 20 and "indnet_lstm_fwd" in protos and "indnet_last_error" in protos
    assert os.path.exists(_lib.LIB_PATH), "build first: python __graft_entry__.py build"
    dll = ctypes.CDLL(_lib.LIB_PATH)
    for name in protos:
        assert hasattr(dll, name), "libindnet.so does not export %s" % name
    assert _lib.LIB.load().indnet_abi_version() == 6


def test_argument_errors_are_reported_not_thrown():
    dll = _lib.LIB.load()
    # bad sizes are rejected before any CUDA call
    rc = dll.indnet_routing_fwd(None, None, None, None, None, 4, 33, 16, 3)
    assert rc != 0 and "warp-per-class limit" in _lib.LIB.last_error()
    rc = dll.indnet_lstm_fwd(None, None, None, None, None, None, None, None, 2, 0, 8, 1.0, 0, None)
    assert rc != 0 and "bad sizes" in _lib.LIB.last_error()
    with pytest.raises(_lib.IndnetError, match="keep_prob"):
        _lib.LIB.call("indnet_lstm_fwd", None, None, None, None, None, None, None, None, 2, 3, 8, 0.0, 0, None)
    # empty inputs are a no-op success
    assert dll.indnet_routing_fwd(None, None, None, None, None, 0, 3, 16, 3) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from induction_b200.engine import ParamStore, tf_param_shapes
    from induction_b200.model import layer_module as lm
    with pytest.raises(_lib.IndnetError, match="no CPU fallback"):
        ParamStore(tf_param_shapes(3, 5, 6, 5, 10))
    with pytest.raises(_lib.IndnetError, match="no CPU fallback"):
        lm.squash(torch.zeros(3, 4))
    with pytest.raises(_lib.IndnetError):
        from induction_b200.model.graph import InductionGraph
        InductionGraph(3, 2, 2, vocab_size=10, sequence_length=5, hidden_size=4)
