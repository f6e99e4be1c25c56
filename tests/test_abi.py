"""The C-ABI library: loads, exports every symbol include/espnet_b200_ctc.h declares, and
rejects bad arguments with the documented status codes (no kernel is launched here)."""
import ctypes
import os
import re

from tests._util import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "espnet_b200_ctc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    names = _declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"


def test_python_binding_covers_header():
    from espnet_b200 import _lib

    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_abi_version(lib):
    assert lib.b200_ctc_abi_version() == 2


def test_workspace_size_monotone(lib):
    a = lib.b200_ctc_loss_workspace_bytes(500, 64, 5000, 150)
    b = lib.b200_ctc_loss_workspace_bytes(500, 64, 5000, 300)
    assert 0 < a < b
    assert lib.b200_ctc_loss_workspace_bytes(0, 64, 5000, 150) == 0


def test_invalid_arguments_are_reported(lib):
    buf = ctypes.create_string_buffer(1 << 16)
    base = (ctypes.addressof(buf) + 255) // 256 * 256
    p = ctypes.c_void_p(base)
    args = dict(T=4, B=1, V=8, st=8, sb=32, Lmax=2)

    def fwd(dtype=0, blank=0, ws=p, ws_bytes=1 << 15, Lmax=2, V=8):
        return lib.b200_ctc_loss_forward(p, dtype, args["T"], args["B"], V, args["st"], args["sb"], p, p, p, p,
                                         Lmax, blank, p, None, 1.0, ws, ws_bytes, None)

    assert fwd(dtype=7) == 1 and b"dtype" in lib.b200_last_error()
    assert fwd(blank=9) == 1 and b"blank" in lib.b200_last_error()
    assert fwd(Lmax=5000) == 3 and b"1023" in lib.b200_last_error()
    assert fwd(ws=ctypes.c_void_p(base + 8)) == 2
    assert fwd(ws_bytes=16) == 4 and b"workspace" in lib.b200_last_error()
    assert fwd(ws=None) == 1
