"""ctypes loader for the CPU oracle (oracle/ige_oracle.c).  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; never by the
product package."""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libige_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(HERE, "ige_oracle.c")
    hdr = os.path.join(HERE, "..", "include", "ige_b200.h")
    stale = (not os.path.exists(SO)) or any(
        os.path.getmtime(f) > os.path.getmtime(SO) for f in (src, hdr) if os.path.exists(f))
    if force or stale:
        subprocess.run(["make", "-C", HERE, "-s", "-B"], check=True)
    return SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(SO)
        for name in dir(_lib):
            pass
    return _lib


def call(name, *structs):
    fn = getattr(lib(), "ige_oracle_" + name)
    fn.restype = C.c_int
    rc = fn(*[C.byref(s) for s in structs])
    if rc != 0:
        raise RuntimeError(f"oracle {name} failed rc={rc}")
