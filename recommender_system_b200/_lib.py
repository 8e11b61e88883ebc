"""ctypes binding of libctrhot.so -- the thin layer between the Python drop-ins and the C-ABI.

Signatures are taken from ``include/ctrhot.h`` itself (parsed once at import), so the header
is the single source of truth.  There is NO CPU fallback: if the shared library is missing the
import of any compute entry point fails loudly.
"""
from __future__ import annotations

import ctypes
import os
import re

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
HEADER = os.path.join(ROOT, "include", "ctrhot.h")
# CTRHOT_LIB: developer knob for tuning builds (profiles/build_variants.py); still the same C-ABI, still CUDA only
LIB_PATH = os.environ.get("CTRHOT_LIB") or os.path.join(PKG, "libctrhot.so")

OK, EARG, ESHAPE, ECUDA, EID = 0, -1, -2, -3, -4
IDS_I32, IDS_F32, IDS_PACKED = 0, 1, 2
LAUNCH_OVERLAP = 0x100
ACT_NONE, ACT_RELU, ACT_PRELU, ACT_DICE, ACT_SIGMOID, ACT_TANH = 0, 1, 2, 3, 4, 5
ACT_BY_NAME = {None: ACT_NONE, "linear": ACT_NONE, "relu": ACT_RELU, "prelu": ACT_PRELU, "dice": ACT_DICE,
               "sigmoid": ACT_SIGMOID, "tanh": ACT_TANH}


class DenseDesc(ctypes.Structure):
    """Mirror of ``ctrhot_dense_desc`` (include/ctrhot.h)."""
    _fields_ = [("weight", ctypes.c_void_p), ("bias", ctypes.c_void_p), ("act_param", ctypes.c_void_p),
                ("packed", ctypes.c_void_p), ("in_dim", ctypes.c_int32), ("out_dim", ctypes.c_int32),
                ("act", ctypes.c_int32), ("alpha_rows", ctypes.c_int32), ("dice_eps", ctypes.c_float),
                ("reserved", ctypes.c_int32)]


_SCALARS = {"uint8_t": ctypes.c_uint8, "int8_t": ctypes.c_int8, "double": ctypes.c_double, "uint64_t": ctypes.c_uint64, "int": ctypes.c_int, "int32_t": ctypes.c_int32, "int64_t": ctypes.c_int64, "float": ctypes.c_float}


def _ctype(decl):
    decl = decl.replace("const", " ").strip()
    if "*" in decl:
        base = decl.replace("*", " ").split()[0]
        if base == "ctrhot_dense_desc":
            return ctypes.POINTER(DenseDesc)
        if base == "char":
            return ctypes.c_char_p
        return ctypes.c_void_p
    base = decl.split()[0]
    return _SCALARS[base]


def parse_header(path=HEADER):
    """-> {name: (restype, [argtypes])} for every function declared in the header."""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    src = re.sub(r"^\s*#[^\n]*", " ", src, flags=re.M)
    src = re.sub(r'extern\s+"C"\s*\{', " ", src)
    src = re.sub(r"typedef\s+(struct|enum)\s*\{.*?\}\s*\w+\s*;", " ", src, flags=re.S)
    src = re.sub(r"enum\s*\{.*?\}\s*;", " ", src, flags=re.S)
    out = {}
    for m in re.finditer(r"([\w\s\*]+?)\b(ctrhot_\w+)\s*\(([^;{]*?)\)\s*;", src):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                a = re.sub(r"\b\w+$", "", a) if not a.endswith("*") else a   # drop the parameter name
                argtypes.append(_ctype(a))
        out[name] = (_ctype(ret) if ret != "void" else None, argtypes)
    return out


SIGNATURES = parse_header()
_lib = None


class CtrhotError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CtrhotError("libctrhot.so not found at %s -- run `python -m recommender_system_b200.build` "
                          "(there is no CPU fallback)" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the header declares a symbol the .so lacks
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc, what=""):
    """Translate a ctrhot_status into the Python exception the reference's TF path would raise."""
    if rc == OK:
        return
    lib = load()
    msg = lib.ctrhot_status_str(int(rc)).decode()
    if rc == ECUDA:
        raise CtrhotError("%s: %s: %s" % (what, msg, lib.ctrhot_last_cuda_error().decode()))
    if rc == EID:
        raise IndexError("%s: %s" % (what, msg))
    raise ValueError("%s: %s" % (what, msg))


LAUNCHES = [0]      # C-ABI compute calls issued by this process (each launches >= 1 kernel)


def call(name, *args):
    lib = load()
    rc = getattr(lib, name)(*args)
    LAUNCHES[0] += 1
    check(rc, name)


def query(name, *args):
    """For the *_bytes functions: returns the (non-negative) value, raises on a negative status."""
    lib = load()
    v = getattr(lib, name)(*args)
    if v < 0:
        check(int(v), name)
    return int(v)
