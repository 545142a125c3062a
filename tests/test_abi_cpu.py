"""The C-ABI library builds, loads without a GPU and exports every entry point that include/*.h
declares; the ctypes binding declares a signature for each of them.  No compute calls here."""
import ctypes
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "speakerverification_b200", "libsvb.so")


def declared_symbols():
    names = []
    for h in sorted(glob.glob(os.path.join(ROOT, "include", "*.h"))):
        src = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names += re.findall(r"\b(svb_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        import __graft_entry__ as g
        g.build()
    return ctypes.CDLL(LIB)


def test_header_declares_the_path():
    names = declared_symbols()
    for must in ("svb_fbank_forward", "svb_ecapa_create", "svb_ecapa_forward", "svb_cosine_trials",
                 "svb_asnorm_stats", "svb_asnorm_trials", "svb_last_error", "svb_version"):
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, "declared in include/*.h but not exported by libsvb.so: %s" % missing
    assert lib.svb_version() == 100


def test_binding_covers_the_header(lib):
    from speakerverification_b200 import _native as nv
    L = nv.lib()
    unbound = [n for n in declared_symbols() if getattr(L, n).argtypes is None and n not in
               ("svb_last_error", "svb_launch_count", "svb_version", "svb_reset_launch_count")]
    assert not unbound, "no ctypes signature for: %s" % unbound


def test_error_convention_without_gpu(lib):
    """Negative return + thread-local message; NULL arguments are refused before any CUDA work."""
    lib.svb_last_error.restype = ctypes.c_char_p
    lib.svb_tdnn_layer.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    rc = lib.svb_tdnn_layer(None, None)
    assert rc < 0
    assert len(lib.svb_last_error()) > 0


def test_product_does_not_import_the_oracle():
    """oracle/ is test infrastructure: no product module may reference it."""
    for py in glob.glob(os.path.join(ROOT, "speakerverification_b200", "**", "*.py"), recursive=True):
        src = open(py).read()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), py
