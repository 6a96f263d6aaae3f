"""ctypes binding of liblgk.so (the C ABI declared in include/lgk.h).

The prototypes are parsed from the header itself, so the Python side can never
drift from the C declarations.  There is no fallback of any kind: if the shared
library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes
import os
import re
from functools import lru_cache

_HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(_HERE), "include", "lgk.h")
LIBRARY = os.path.join(_HERE, "liblgk.so")

F32, BF16 = 0, 1
SUM, MEAN, MAX = 0, 1, 2
SELF_NONE, SELF_REPLACE, SELF_ADD = 0, 1, 2
ACT_NONE, ACT_ELU, ACT_RELU = 0, 1, 2

_SCALARS = {
    "int": ctypes.c_int,
    "int64_t": ctypes.c_int64,
    "uint64_t": ctypes.c_uint64,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
}


def parse_header(path: str = HEADER) -> dict[str, tuple]:
    """{name: (restype, [argtypes], [argnames])} for every LGK_API declaration."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    protos = {}
    for ret, name, args in re.findall(r"LGK_API\s+([\w\s\*]+?)\s*\b(lgk_\w+)\s*\(([^)]*)\)\s*;", text, flags=re.S):
        ret = " ".join(ret.split())
        if ret == "void":
            restype = None
        elif ret == "const char*" or ret == "const char *":
            restype = ctypes.c_char_p
        else:
            restype = _SCALARS[ret]
        argtypes, argnames = [], []
        args = " ".join(args.split())
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                    argnames.append(a.split("*")[-1].strip())
                else:
                    toks = [t for t in a.split() if t != "const"]
                    argtypes.append(_SCALARS[toks[0]])
                    argnames.append(toks[-1])
        protos[name] = (restype, argtypes, argnames)
    return protos


class _Lib:
    def __init__(self):
        if not os.path.exists(LIBRARY):
            raise RuntimeError(
                f"{LIBRARY} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C lesion_gnn_b200/csrc`). lesion_gnn_b200 has no CPU or PyTorch fallback."
            )
        self._dll = ctypes.CDLL(LIBRARY)
        self.protos = parse_header()
        for name, (restype, argtypes, _) in self.protos.items():
            fn = getattr(self._dll, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = restype
            fn.argtypes = argtypes
        self._dll.lgk_last_error.restype = ctypes.c_char_p
        # optional observer hook(phase, name, args) around every call: bench.py uses it to bracket each entry
        # point with CUDA events for the per-kernel roofline; None in normal operation
        self.hook = None

    def last_error(self) -> str:
        return self._dll.lgk_last_error().decode()

    def call(self, name: str, *args):
        """Call a status-returning entry point; raise RuntimeError(lgk_last_error()) on failure."""
        hook = self.hook
        if hook is not None:
            hook("begin", name, args)
        rc = getattr(self._dll, name)(*args)
        if rc != 0:
            raise RuntimeError(f"{name} failed ({rc}): {self.last_error()}")
        if hook is not None:
            hook("end", name, args)

    def query(self, name: str, *args):
        """Call a value-returning entry point (workspace sizes, counters)."""
        return getattr(self._dll, name)(*args)


@lru_cache(maxsize=None)
def lib() -> _Lib:
    return _Lib()
