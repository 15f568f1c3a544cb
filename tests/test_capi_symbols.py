"""The C-ABI shared library loads and exports every symbol include/synthpop_cuda.h declares (no GPU needed)."""
import ctypes as C
import os

from synthpop import lib as splib


def test_library_exports_every_declared_symbol():
    assert os.path.exists(splib.LIB_PATH), "libsynthpop_cuda.so is not built (run __graft_entry__.build())"
    dll = C.CDLL(splib.LIB_PATH)
    names = splib.declared_symbols()
    assert len(names) >= 15
    missing = [n for n in names if not hasattr(dll, n)]
    assert not missing, missing


def test_binding_declares_prototypes():
    lib = splib.load()
    assert lib.sp_version().startswith(b"synthpop-cuda")
