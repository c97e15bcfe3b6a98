"""The C-ABI library loads on a CPU-only box and exports every symbol include/nosplan.h declares
(no compute calls here: there is no GPU and no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    return ctypes.CDLL(g.LIB)


def declared_functions():
    src = open(os.path.join(ROOT, "include", "nosplan.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nosplan_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_exported(lib):
    names = declared_functions()
    assert "nosplan_plan" in names and "nosplan_create" in names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/nosplan.h but not exported"


def test_python_binding_lists_the_same_symbols():
    from nos_b200 import capi
    assert sorted(capi.EXPORTED_SYMBOLS) == declared_functions()


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nos_b200 import capi
    with pytest.raises(capi.NosplanError):
        capi.Engine(0)


def test_struct_sizes_match_header(lib):
    """ctypes mirrors of the C structs: compile a tiny probe with gcc and compare sizeof."""
    import subprocess
    import tempfile
    from nos_b200 import capi
    probe = r'''
#include <stdio.h>
#include "nosplan.h"
int main(void){printf("%zu %zu %zu %zu\n", sizeof(nosplan_snapshot), sizeof(nosplan_pods), sizeof(nosplan_partitioning), sizeof(nosplan_stats));return 0;}
'''
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "p.c")
        open(c, "w").write(probe)
        exe = os.path.join(d, "p")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        sizes = [int(x) for x in subprocess.check_output([exe]).split()]
    assert sizes == [ctypes.sizeof(capi.CSnapshot), ctypes.sizeof(capi.CPods), ctypes.sizeof(capi.CPartitioning),
                     ctypes.sizeof(capi.CStats)]
