"""The C-ABI boundary: the product library loads without a GPU and exports every function include/bwamem_b200.h declares
(no compute call is made here), and the Python binding's own symbol list is a subset of the header."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "bwamem_b200.h")
PRODUCT = os.path.join(ROOT, "bwa-mem-gpu_b200", "libbwamem_b200.so")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)                 # comments out
    text = re.sub(r"//[^\n]*", " ", text)
    text = re.sub(r"^\s*#[^\n]*(\\\n[^\n]*)*", " ", text, flags=re.M)   # preprocessor lines (with continuations)
    names = set()
    for stmt in text.split(";"):
        if "typedef" in stmt or "(" not in stmt:
            continue
        head = stmt[: stmt.index("(")]
        if "{" in head or "}" in head and "extern" not in head:
            head = head.split("}")[-1]
        m = re.search(r"([A-Za-z_]\w*)\s*$", head)
        if m and re.match(r"(b200_|mem_|bwt_|ksw_)", m.group(1)):
            names.add(m.group(1))
    return sorted(names)


def test_header_declares_functions():
    names = declared_functions()
    assert len(names) >= 50 and "b200_reads_to_sam" in names and "mem_process_seqs" in names and "b200_ksw_extend2_noalloc" in names


@pytest.mark.skipif(not os.path.exists(PRODUCT), reason="product library not built (python -c 'import __graft_entry__ as g; g.build()')")
def test_product_library_exports_every_declared_function():
    lib = C.CDLL(PRODUCT, mode=os.RTLD_LOCAL)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_symbols_are_declared():
    import sys
    sys.path.insert(0, ROOT)
    from bwa_mem_gpu_b200 import lib as L
    internal = {"b200_microbench", "b200_batch_counters", "b200_hostidx_build", "b200_hostidx_free", "b200_hostidx_info", "b200_hostidx_bwt",
                "b200_hostidx_sa", "b200_hostidx_pac"}
    decl = set(declared_functions())
    undeclared = [s for s in L.exported_symbols() if s not in decl and s not in internal]
    assert not undeclared, undeclared
