"""The C-ABI library loads and exports every symbol include/gsb.h declares (no GPU needed)."""
import ctypes
import os

from gslam_b200 import _lib


def test_library_exists():
    assert os.path.exists(_lib.LIB_PATH), "run `make` / __graft_entry__.build() first"


def test_every_header_symbol_is_exported_and_bound():
    names = _lib.header_symbols()
    assert len(names) >= 14
    L = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/gsb.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in gslam_b200/_lib.py"
    assert set(_lib.SIGNATURES) == set(names)


def test_version_and_error_string():
    L = _lib.lib()
    assert L.gsb_version() >= 100
    assert isinstance(L.gsb_last_error(), bytes)


def test_bad_arguments_return_codes_not_crashes():
    L = _lib.lib()
    # width <= 0 -> GSB_EINVAL before any CUDA call
    rc = L.gsb_project_fwd(None, None, None, None, None, None, None, None, 1, 1, 0, 0,
                           0.3, 0.01, 1e10, 0.0, 1, 1, None, None, None, None, None, None, -1, 0, None)
    assert rc == -1
    assert b"bad argument" in L.gsb_last_error()
    # empty problems are a no-op success
    assert L.gsb_project_fwd(None, None, None, None, None, None, None, None, 0, 0, 8, 8,
                             0.3, 0.01, 1e10, 0.0, 1, 1, None, None, None, None, None, None, -1, 0, None) == 0


def test_no_oracle_import_in_product():
    pkg = os.path.dirname(_lib.__file__)
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "import oracle" not in src and "from oracle" not in src, fn
