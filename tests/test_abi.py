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


def test_header_is_plain_c_and_a_c_program_links_against_the_library(dx, tmp_path):
    """include/dxgpu.h must compile as C99 and as C++ on its own (a cgo / JNI / ctypes binding sees nothing else), and a C program
    built against it and libdxgpu.so runs the entry points that need no GPU (tests/native/abi_smoke.c)."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = os.path.join(root, "include", "dxgpu.h")
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", hdr], check=True)
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-fsyntax-only", "-x", "c++", hdr], check=True)
    libdir = os.path.dirname(dx.LIB_PATH)
    exe = str(tmp_path / "abi_smoke")
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I" + os.path.join(root, "include"),
                    os.path.join(root, "tests", "native", "abi_smoke.c"), "-o", exe, "-L" + libdir, "-l:" + os.path.basename(dx.LIB_PATH),
                    "-Wl,-rpath," + libdir], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    assert "SELECT a, COUNT(*) AS n FROM t WHERE (b > 1) GROUP BY a" in out and "T1\tQuery\tSELECT * FROM x" in out
