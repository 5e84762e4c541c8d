"""ctypes binding of ``csrc/libindnet.so`` generated from ``include/indnet.h``.

The prototypes are parsed out of the public header, so the Python side cannot drift from the C-ABI the
header declares.  Loading fails loudly: there is no CPU fallback behind these calls.
"""
from __future__ import annotations

import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.normpath(os.path.join(_HERE, "..", "include", "indnet.h"))
LIB_PATH = os.path.join(_HERE, "csrc", "libindnet.so")

_SCALARS = {
    "int": ctypes.c_int,
    "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
    "indnet_stream": ctypes.c_void_p,
}


def parse_header(path: str = HEADER):
    """Returns {name: (restype, [(ctype, param_name), ...])} for every prototype in the header."""
    with open(path) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    protos = {}
    for m in re.finditer(r"(const\s+char\s*\*|long long|int)\s+(indnet_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        ret, name, params = m.group(1), m.group(2), m.group(3)
        restype = ctypes.c_char_p if "char" in ret else (ctypes.c_longlong if "long" in ret else ctypes.c_int)
        args = []
        params = " ".join(params.split())
        if params and params != "void":
            for p in params.split(","):
                p = p.strip()
                if "*" in p:
                    args.append((ctypes.c_void_p, p.split("*")[-1].strip()))
                else:
                    toks = p.split()
                    ctype_name, pname = " ".join(toks[:-1]), toks[-1]
                    ctype_name = ctype_name.replace("const ", "").strip()
                    args.append((_SCALARS[ctype_name], pname))
        protos[name] = (restype, args)
    return protos


class IndnetError(RuntimeError):
    pass


class _Lib:
    def __init__(self):
        self._dll = None
        self.protos = parse_header()

    def load(self):
        if self._dll is not None:
            return self._dll
        if not os.path.exists(LIB_PATH):
            raise IndnetError(
                "libindnet.so is not built (%s). Run `python __graft_entry__.py build` or "
                "`python induction-network-on-fewrel_b200/csrc/build.py`; there is no CPU fallback." % LIB_PATH)
        dll = ctypes.CDLL(LIB_PATH)
        for name, (restype, args) in self.protos.items():
            fn = getattr(dll, name)   # AttributeError here == header/library mismatch: let it surface
            fn.restype = restype
            fn.argtypes = [a[0] for a in args]
        self._dll = dll
        return dll

    def last_error(self) -> str:
        return self.load().indnet_last_error().decode("utf-8", "replace")

    def launch_count(self) -> int:
        return int(self.load().indnet_launch_count())

    def call(self, name: str, *args):
        fn = getattr(self.load(), name)
        rc = fn(*args)
        if rc != 0:
            raise IndnetError("%s failed (code %d): %s" % (name, rc, self.last_error()))


LIB = _Lib()


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise IndnetError("induction_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback.")


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def current_stream():
    import torch
    return torch.cuda.current_stream().cuda_stream
