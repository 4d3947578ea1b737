"""ctypes binding of libdtf_b200_cuda.so. Signatures are parsed from include/dtf_b200.h (single source of truth).

The library is loaded lazily; on a machine with a GPU a missing library is a hard error (no silent fallback).
"""
import ctypes
import os
import re

import torch

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_HEADER = os.path.join(os.path.dirname(_PKG), "include", "dtf_b200.h")
_LIB_PATH = os.path.join(_PKG, "lib", "libdtf_b200_cuda.so")

_SCALARS = {
    "int": ctypes.c_int,
    "long long": ctypes.c_longlong,
    "unsigned int": ctypes.c_uint,
    "unsigned long long": ctypes.c_ulonglong,
    "size_t": ctypes.c_size_t,
    "float": ctypes.c_float,
}
ERRORS = {1: "bad argument", 2: "misaligned pointer / size", 3: "CUDA error", 4: "driver entry point missing",
          5: "tensor map encode failed", 6: "size beyond supported maximum", 7: "unsupported configuration"}


class DtfError(RuntimeError):
    pass


def parse_header(path=_HEADER):
    """Return {name: (restype, [argtypes], [argnames])} for every declaration in the public header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    text = re.sub(r"#[^\n]*", "", text)
    out = {}
    for m in re.finditer(r"(int|void\s*\*)\s+(dtf_\w+)\s*\(([^)]*)\)\s*;", text):
        rtype, name, args = m.group(1), m.group(2), m.group(3).strip()
        argtypes, argnames = [], []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                    argnames.append(a.split("*")[-1].strip())
                else:
                    ty, nm = a.rsplit(" ", 1)
                    ty = ty.replace("const ", "").strip()
                    argtypes.append(_SCALARS[ty])
                    argnames.append(nm)
        out[name] = (ctypes.c_int if rtype == "int" else ctypes.c_void_p, argtypes, argnames)
    return out


_lib = None
_sigs = None


def available():
    return os.path.exists(_LIB_PATH)


def lib():
    global _lib, _sigs
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise DtfError("native library %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(no silent fallback on the CUDA path)" % _LIB_PATH)
        _lib = ctypes.CDLL(_LIB_PATH)
        _sigs = parse_header()
        for name, (res, argtypes, _) in _sigs.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = argtypes
    return _lib


def _conv(x):
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        return x.data_ptr()
    return x


launch_count = 0


def call(name, *args):
    """Call an int-returning entry point; tensors become device pointers; non-zero return raises."""
    global launch_count
    fn = getattr(lib(), name)
    rc = fn(*[_conv(a) for a in args])
    launch_count += 1
    if rc != 0:
        raise DtfError("%s failed: %s (code %d)" % (name, ERRORS.get(rc, "?"), rc))
    return rc


def call_raw(name, *args):
    fn = getattr(lib(), name)
    return fn(*[_conv(a) for a in args])


def stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream
