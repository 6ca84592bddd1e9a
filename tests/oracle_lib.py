"""ctypes loader for the CPU oracle (oracle/build/liboracle.so). Test infrastructure only."""
import ctypes
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(ROOT, "oracle", "build", "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_call.restype = ctypes.c_void_p
        _LIB.oracle_call.argtypes = [ctypes.c_char_p]
        _LIB.oracle_free.argtypes = [ctypes.c_void_p]
    return _LIB


def call(req):
    l = lib()
    p = l.oracle_call(json.dumps(req).encode())
    try:
        out = ctypes.string_at(p).decode()
    finally:
        l.oracle_free(p)
    return json.loads(out)
