"""CPU: libnbrm.so loads and exports every entry point include/nbrm.h declares.
No compute call is made here (the container has no GPU)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols(header="nbrm.h"):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nbrm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    import __graft_entry__ as graft
    if not os.path.exists(pkg.lib_path()):
        graft.build()
    lib = ctypes.CDLL(pkg.lib_path())
    syms = declared_symbols()
    assert len(syms) >= 15
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_host_library_exports_every_declared_symbol(pkg):
    import __graft_entry__ as graft
    host = os.path.join(os.path.dirname(pkg.lib_path()), "libnbrm_host.so")
    if not os.path.exists(host):
        graft.build()
    ctypes.CDLL(pkg.lib_path(), mode=ctypes.RTLD_GLOBAL)
    lib = ctypes.CDLL(host)
    syms = [s for s in declared_symbols("nbrm_host.h") if s.startswith(("nbrm_store", "nbrm_reconciler", "nbrm_host"))]
    assert len(syms) >= 10
    assert not [s for s in syms if not hasattr(lib, s)]


def test_no_cpu_fallback_without_device(pkg):
    """On a box without a GPU the product path must fail loudly, not fall back."""
    import pytest
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    with pytest.raises(pkg.NbrmError):
        pkg.Context(0)


def test_product_never_imports_oracle():
    """The package and the CUDA sources never reference oracle/ or tests/hostplan."""
    pkgdir = os.path.join(ROOT, "kubeflow-rm_b200")
    for dirpath, _, files in os.walk(pkgdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inc", ".cc")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "hostplan" not in text, f
