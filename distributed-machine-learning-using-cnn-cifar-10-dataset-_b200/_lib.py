"""ctypes binding of libb200cnn.so (include/b200cnn.h).  There is no CPU fallback: if the
library cannot be built/loaded, or no CUDA device is present when a compute call is made,
the caller gets an exception."""
from __future__ import annotations

import ctypes as C
import os
import re

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(HERE, "..", "include", "b200cnn.h")

OK, E_INVALID, E_CUDA, E_STATE, E_NODEVICE = 0, -1, -2, -3, -4
MODE_FP32, MODE_BF16 = 0, 1
NUM_PARAMS = 10


class B200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200cnn error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "batch", "img_h", "img_w", "crop_h", "crop_w", "c_in", "c1", "c2", "f1", "f2", "n_cls", "mode")]


_CTYPE = {
    "void*": C.c_void_p, "const void*": C.c_void_p, "float*": C.c_void_p, "const float*": C.c_void_p,
    "uint8_t*": C.c_void_p, "const uint8_t*": C.c_void_p, "int32_t*": C.c_void_p,
    "const int32_t*": C.c_void_p, "int32_t": C.c_int32, "int64_t": C.c_int64, "float": C.c_float,
    "int": C.c_int, "const char*": C.c_char_p, "b200_handle": C.c_void_p,
    "int64_t*": C.POINTER(C.c_int64), "b200_handle*": C.POINTER(C.c_void_p),
    "const b200_config*": C.POINTER(Config), "float**": C.POINTER(C.c_void_p),
    "int32_t**": C.POINTER(C.c_void_p), "void**": C.POINTER(C.c_void_p),
    "const void* const*": C.POINTER(C.c_void_p), "void* const*": C.POINTER(C.c_void_p), "uint32_t*": C.c_void_p,
    "const uint8_t* const*": C.POINTER(C.c_void_p), "uint64_t*": C.c_void_p, "uint64_t": C.c_uint64,
}
_PTR_OUT = {"b200_profile_step": {5: C.POINTER(C.c_float), 7: C.POINTER(C.c_int32)},
            "b200_launches_per_step": {1: C.POINTER(C.c_int32)},
            "b200_tensor_ptr": {3: C.POINTER(C.c_int64), 4: C.POINTER(C.c_int32)}}


def declared_functions(header=HEADER):
    """[(return type, name, [arg types])] parsed from the header -- the single source of truth
    for the binding, and what tests/test_abi.py checks the .so against."""
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = []
    for m in re.finditer(r"(?m)^(const char\*|int)\s+(b200_\w+)\s*\(([^)]*)\)\s*;", src):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        types = []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                t = re.sub(r"\s*\w+$", "", a) if not a.endswith("*") else a   # drop the parameter name
                t = t.replace(" *", "*").strip()
                types.append(t)
        out.append((ret, name, types))
    return out


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("B200CNN_LIB") or _build.build()      # (diagnostics: a B200_TRACE=1 build kept beside the product library)
    lib = C.CDLL(path)
    for ret, name, types in declared_functions():
        fn = getattr(lib, name)
        fn.restype = C.c_char_p if ret == "const char*" else C.c_int
        fn.argtypes = [_PTR_OUT.get(name, {}).get(i, _CTYPE[t]) for i, t in enumerate(types)]
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise B200Error(rc, load().b200_last_error().decode())
    return rc
