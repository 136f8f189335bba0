"""ctypes loader for the CPU oracle (oracle/_build/liboracle.so). Test infrastructure only.

Prototypes are parsed from oracle/*.cc (`ORACLE_API <ret> name(args)`), so the binding
follows the source. numpy arrays are passed as pointers; the caller owns dtype/shape."""
import ctypes
import re
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
SO = ROOT / "oracle" / "_build" / "liboracle.so"

_SCALARS = {
    "int": ctypes.c_int, "float": ctypes.c_float, "double": ctypes.c_double,
    "size_t": ctypes.c_size_t, "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong, "uint64_t": ctypes.c_uint64,
    "int64_t": ctypes.c_int64, "unsigned": ctypes.c_uint,
}
_RE = re.compile(r"ORACLE_API\s+(void|int|double|float|size_t)\s+(oracle_[a-z0-9_]+)\s*\(([^)]*)\)", re.S)


def build():
    subprocess.run(["make", "-s", "-C", str(ROOT / "oracle")], check=True)


class Oracle:
    def __init__(self):
        if not SO.exists():
            build()
        self.cdll = ctypes.CDLL(str(SO))
        for src in sorted((ROOT / "oracle").glob("*.cc")):
            text = re.sub(r"//[^\n]*", "", src.read_text())
            for m in _RE.finditer(text):
                ret, name, args = m.groups()
                argtypes = []
                for a in args.split(","):
                    a = " ".join(a.split())
                    if not a or a == "void":
                        continue
                    if "*" in a:
                        argtypes.append(ctypes.c_void_p)
                    else:
                        base = " ".join(a.replace("const ", "").split(" ")[:-1])
                        argtypes.append(_SCALARS[base])
                fn = getattr(self.cdll, name)
                fn.argtypes = argtypes
                fn.restype = {"void": None, "int": ctypes.c_int, "double": ctypes.c_double,
                              "float": ctypes.c_float, "size_t": ctypes.c_size_t}[ret]

    def __getattr__(self, name):
        fn = getattr(self.cdll, "oracle_" + name)

        def call(*args):
            conv = []
            for a in args:
                if isinstance(a, np.ndarray):
                    assert a.flags["C_CONTIGUOUS"], "oracle operands must be C-contiguous"
                    conv.append(a.ctypes.data)
                elif a is None:
                    conv.append(None)
                elif isinstance(a, (list, tuple)):
                    arr = (ctypes.c_longlong * len(a))(*a)
                    conv.append(ctypes.cast(arr, ctypes.c_void_p))
                else:
                    conv.append(a)
            return fn(*conv)

        return call


_o = None


def load():
    global _o
    if _o is None:
        _o = Oracle()
    return _o
