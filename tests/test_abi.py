"""The C-ABI library loads and exports every symbol include/dxgpu.h declares (no compute calls)."""
import ctypes
import os


def test_library_exports_every_declared_symbol(dx):
    assert os.path.exists(dx.LIB_PATH), "libdxgpu.so not built (run __graft_entry__.build())"
    L = ctypes.CDLL(dx.LIB_PATH)
    names = dx.declared_symbols()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, f"declared in include/dxgpu.h but not exported: {missing}"


def test_no_oracle_in_product_path():
    """The product package must never import or link the oracle."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "data-accelerator_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "libdxo" not in text and "import dxo" not in text and "oracle/" not in text.replace(
                    "oracle/dxo_core.c", "").replace("oracle/dxo.py", ""), f
