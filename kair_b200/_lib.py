"""ctypes binding of the C-ABI library, generated from include/kair_b200.h.

The header is the single source of truth: argument structs and function prototypes are parsed from it, so the
Python side cannot drift from the C side.  The library itself is built in-tree by kair_b200/build.py (nvcc,
sm_100a).  There is no fallback: a missing library or a failing call raises RuntimeError.
"""
import ctypes
import os
import re
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkair_b200.so")
HEADER_PATH = os.path.join(_HERE, "..", "include", "kair_b200.h")

_SCALARS = {
    "int32_t": ctypes.c_int32, "int64_t": ctypes.c_int64, "float": ctypes.c_float, "int": ctypes.c_int,
    "double": ctypes.c_double, "uint32_t": ctypes.c_uint32,
}


def _strip_comments(src):
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    return re.sub(r"//[^\n]*", " ", src)


def _parse_header(path):
    src = _strip_comments(open(path).read())
    structs = {}
    for m in re.finditer(r"typedef\s+struct\s+(\w+)\s*\{(.*?)\}\s*(\w+)\s*;", src, flags=re.S):
        name, body = m.group(3), m.group(2)
        fields = []
        for decl in body.split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            if "*" in decl:
                # every pointer flavour is passed as an opaque address
                fname = decl.rsplit("*", 1)[1].strip()
                fields.append((fname, ctypes.c_void_p))
            else:
                tname, rest = decl.split(" ", 1)
                for fname in rest.split(","):
                    fields.append((fname.strip(), _SCALARS[tname]))
        structs[name] = type(name, (ctypes.Structure,), {"_fields_": fields})
    protos = {}
    for m in re.finditer(r"(const\s+char\s*\*|int)\s+(kair_\w+)\s*\(([^)]*)\)\s*;", src):
        ret = ctypes.c_char_p if "char" in m.group(1) else ctypes.c_int
        args = []
        params = m.group(3).strip()
        if params and params != "void":
            for prm in params.split(","):
                prm = " ".join(prm.split())
                if "*" in prm:
                    base = prm.split("*")[0].replace("const", "").strip()
                    args.append(ctypes.POINTER(structs[base]) if base in structs else ctypes.c_void_p)
                else:
                    args.append(_SCALARS[prm.split(" ")[0]])
        protos[m.group(2)] = (ret, args)
    return structs, protos


STRUCTS, SIGNATURES = _parse_header(HEADER_PATH)

_lib = None
_lock = threading.Lock()


def load():
    """Load (once) and return the ctypes handle; raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "kair_b200: %s not found. Build it with `python -m kair_b200.build` (needs nvcc); "
                "there is no CPU / PyTorch fallback." % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc, what):
    if rc != 0:
        msg = load().kair_last_error()
        raise RuntimeError("kair_b200.%s failed (rc=%d): %s" % (what, rc, msg.decode() if msg else "?"))


def call(name, args_struct, stream):
    """Invoke `int name(const args*, void* stream)` and raise on error."""
    check(getattr(load(), name)(ctypes.byref(args_struct), stream), name)
