"""induction-network-on-fewrel_b200: B200-native (sm_100a) episode pipeline of the Induction Network.

Host side mirrors the reference's Python API (model/layer_module.py, model/graph.py,
util/data_loader.py); every hot op underneath is a hand-written CUDA kernel reached through the
C-ABI library ``csrc/libindnet.so`` (declared in ``include/indnet.h``).  There is no CPU fallback:
compute entry points raise if the library or a CUDA device is missing.
"""
__version__ = "0.1.0"
