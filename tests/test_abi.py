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


def test_handle_struct_layout_matches_header(tmp_path):
    """The ctypes mirrors in _clib.py must have the sizes a C compiler gives the structs of include/xgnn_b200.h."""
    import subprocess
    from pathlib import Path
    inc = Path(__file__).resolve().parent.parent / "include"
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "xgnn_b200.h"\n#include "samgraph_abi.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu\\n", sizeof(xg_graph), sizeof(xg_feature_store), '
                   'sizeof(xg_sampler_config), sizeof(xg_batch_view), sizeof(samgraph_tensor_view));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c99", "-I", str(inc), str(src), "-o", str(exe)], check=True)  # headers are plain C
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    from xgnn_b200.samgraph.torch import _TensorView
    assert sizes == [ctypes.sizeof(_clib.XgGraph), ctypes.sizeof(_clib.XgFeatureStore),
                     ctypes.sizeof(_clib.XgSamplerConfig), ctypes.sizeof(_clib.XgBatchView), ctypes.sizeof(_TensorView)]


def test_invalid_arguments_are_reported_not_crashed():
    rc = _clib.lib.xg_hashtable_reset(None, None)
    assert rc == -2 and b"NULL" in _clib.lib.xg_last_error()
