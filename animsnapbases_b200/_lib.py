"""ctypes binding of libasb.so.  Prototypes are parsed from include/asb.h so the Python side can never drift
from the C-ABI.  There is no CPU fallback: a missing library or a missing GPU is a hard error."""
from __future__ import annotations

import ctypes
import os
import re

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
HEADER = os.path.join(_ROOT, "include", "asb.h")
LIBPATH = os.path.join(_PKG, "lib", "libasb.so")

_SCALARS = {"int": ctypes.c_int, "int64_t": ctypes.c_int64, "double": ctypes.c_double, "void": None,
            "int32_t": ctypes.c_int32}


class AsbError(RuntimeError):
    pass


def parse_header(path: str = HEADER):
    """-> {name: (restype, [argtypes], [argnames])} for every `asb_*` function declared in the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(asb_\w+)\s*\(([^;{}]*?)\)\s*;", text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        if "typedef" in ret:
            continue

        def ctype(decl: str):
            decl = decl.strip()
            if "*" in decl or decl.startswith("asb_stream_t"):
                if decl.replace(" ", "").startswith("constchar*"):
                    return ctypes.c_char_p
                return ctypes.c_void_p
            base = decl.replace("const", "").split()[0]
            return _SCALARS[base]

        argtypes, argnames = [], []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                argtypes.append(ctype(a))
                argnames.append(re.split(r"[\s\*]+", a)[-1])
        protos[name] = (ctype(ret) if ret != "void" else None, argtypes, argnames)
    return protos


_lib = None
_protos = None


def load():
    """Load libasb.so and attach prototypes.  Raises AsbError when the library has not been built."""
    global _lib, _protos
    if _lib is not None:
        return _lib
    if not os.path.exists(LIBPATH):
        raise AsbError(f"{LIBPATH} is missing: run `python -m animsnapbases_b200.build` "
                       "(there is no CPU fallback for the basis engine)")
    lib = ctypes.CDLL(LIBPATH, mode=ctypes.RTLD_GLOBAL)
    _protos = parse_header()
    for name, (res, argtypes, _) in _protos.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export what the header declares
        fn.restype = res
        fn.argtypes = argtypes
    _lib = lib
    return lib


def prototypes():
    return parse_header() if _protos is None else _protos


def check(status: int, what: str = ""):
    if status != 0:
        msg = load().asb_last_error()
        raise AsbError(f"{what} failed with status {status}: {msg.decode() if msg else ''}")
