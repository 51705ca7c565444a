"""CPU: the C-ABI library loads and exports every symbol the headers under include/ declare.
No compute call is made (there is no GPU here)."""
import ctypes

from xgnn_b200 import _clib


def test_library_loads_and_exports_every_declared_symbol():
    names = _clib.declared_symbols()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(_clib.lib, n)]
    assert not missing, f"declared in include/*.h but not exported: {missing}"


def test_version_and_error_string():
    assert _clib.lib.xg_version() >= 100
    assert isinstance(_clib.lib.xg_last_error(), bytes)
    assert _clib.lib.xg_launch_count() == 0 or _clib.lib.xg_launch_count() > 0


def test_handle_struct_layout_matches_header():
    # xg_graph: 2 * 17 pointers + 4 u32; xg_feature_store: 16 + 2 pointers + 2 u32 + u64
    assert ctypes.sizeof(_clib.XgGraph) == 2 * 17 * 8 + 16
    assert ctypes.sizeof(_clib.XgFeatureStore) == 18 * 8 + 8 + 8


def test_invalid_arguments_are_reported_not_crashed():
    rc = _clib.lib.xg_hashtable_reset(None, None)
    assert rc == -2 and b"NULL" in _clib.lib.xg_last_error()
