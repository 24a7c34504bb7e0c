"""CPU-side checks of the C-ABI boundary: the library loads and exports every symbol include/hoomd_b200.h declares."""
import ctypes
import os
import re

from hoomd_b200 import _lib


def _header_symbols():
    text = open(_lib.HEADER_PATH).read()
    return sorted(set(re.findall(r"B200_API\s+[^;(]*?\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_list_agree():
    assert _header_symbols() == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build the extension first: python -c 'import __graft_entry__ as g; g.build()'"
    L = ctypes.CDLL(_lib.LIB_PATH)
    missing = [s for s in _header_symbols() if not hasattr(L, s)]
    assert not missing, missing


def test_host_only_entry_points():
    """Entry points that are pure host arithmetic run without a GPU: BoxDim construction, cell dimensions, seed hash."""
    L = _lib.lib()
    b = _lib.make_box(68.399, 68.399, 68.399)
    assert abs(b.hi[0] - 34.1995) < 1e-4 and b.lo[0] == -b.hi[0] and tuple(b.periodic) == (1, 1, 1)
    dim = _lib.Dim3()
    gw = (ctypes.c_float * 3)()
    L.b200_cell_list_dimensions(ctypes.byref(b), 3.4, 1, 0.0, ctypes.byref(dim), gw)
    assert (dim.x, dim.y, dim.z) == (20, 20, 20)  # floor(68.399 / 3.4) = 20
    # TwoStepLangevinBase seed hash (hoomd/md/TwoStepLangevinBase.cc:54)
    s = 42
    h = (s * 0x12345677 + 0x12345) & 0xffffffff
    h ^= h >> 16
    h = (h * 0x45679) & 0xffffffff
    assert L.b200_langevin_hash_seed(42) == h
