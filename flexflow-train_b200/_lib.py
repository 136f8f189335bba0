"""ctypes binding of libffk.so, generated from the prototypes in include/*.h.

The header files are the single statement of the C ABI; this module parses them so
that the Python-side argument types can never drift from the declarations. There is
no fallback: if the shared library is missing, importing the ops fails loudly.
"""
from __future__ import annotations

import ctypes
import os
import re
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
INCLUDE_DIR = REPO_ROOT / "include"
LIB_PATH = PKG_DIR / "libffk.so"

_CTYPES = {
    "int": ctypes.c_int,
    "unsigned": ctypes.c_uint,
    "unsigned int": ctypes.c_uint,
    "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
    "size_t": ctypes.c_size_t,
    "int64_t": ctypes.c_int64,
    "uint64_t": ctypes.c_uint64,
    "int32_t": ctypes.c_int32,
    "uint32_t": ctypes.c_uint32,
}

_PROTO_RE = re.compile(
    r"^\s*(int|size_t|char const \*|void)\s*(ffk_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", re.M | re.S
)


def _strip_comments(text: str) -> str:
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return re.sub(r"//[^\n]*", " ", text)


def _arg_ctype(arg: str):
    arg = " ".join(arg.split())
    if arg in ("void", ""):
        return None
    if "*" in arg:
        return ctypes.c_void_p
    # drop the parameter name
    parts = arg.replace("const ", "").split(" ")
    base = " ".join(parts[:-1]) if len(parts) > 1 else parts[0]
    if base not in _CTYPES:
        raise RuntimeError(f"ffk binding: unknown C type in header: {arg!r}")
    return _CTYPES[base]


def parse_headers(include_dir: Path = INCLUDE_DIR):
    """Return {name: (restype, [argtypes])} for every prototype in include/*.h."""
    protos = {}
    for h in sorted(include_dir.glob("*.h")):
        text = _strip_comments(h.read_text())
        for m in _PROTO_RE.finditer(text):
            ret, name, args = m.group(1), m.group(2), m.group(3)
            argtypes = [t for t in (_arg_ctype(a) for a in args.split(",")) if t is not None]
            restype = {"int": ctypes.c_int, "size_t": ctypes.c_size_t, "void": None}.get(
                ret, ctypes.c_char_p
            )
            protos[name] = (restype, argtypes)
    return protos


class FFKError(RuntimeError):
    def __init__(self, name: str, code: int, message: str):
        super().__init__(f"{name} failed ({code}): {message}")
        self.code = code


class _Lib:
    def __init__(self):
        if not LIB_PATH.exists():
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (make -C flexflow-train_b200/csrc). There is no CPU fallback."
            )
        self.cdll = ctypes.CDLL(os.fspath(LIB_PATH))
        self.protos = parse_headers()
        self.launches = 0  # number of C-ABI compute calls made through `call`
        for name, (restype, argtypes) in self.protos.items():
            fn = getattr(self.cdll, name)  # AttributeError if a declared symbol is missing
            fn.restype = restype
            fn.argtypes = argtypes

    def last_error(self) -> str:
        return self.cdll.ffk_last_error().decode()

    def call(self, name: str, *args):
        rc = getattr(self.cdll, name)(*args)
        self.launches += 1
        if rc < 0:
            raise FFKError(name, rc, self.last_error())
        return rc


_lib = None


def lib() -> _Lib:
    global _lib
    if _lib is None:
        _lib = _Lib()
    return _lib
