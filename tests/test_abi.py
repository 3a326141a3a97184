"""The C-ABI library loads without a GPU, exports every symbol include/igmk.h declares, and the
ctypes mirrors have the C layout."""
import ctypes as C
import os
import re
import subprocess
import tempfile

from conftest import ROOT
from igm_b200 import _abi


def _header_functions():
    txt = open(os.path.join(ROOT, "include", "igmk.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(igmk_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported_and_bound():
    declared = _header_functions()
    assert len(declared) >= 25
    lib = C.CDLL(_abi.lib_path())
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == _abi.declared_symbols()


def test_struct_layouts_match_the_header():
    src = r'''
    #include <stdio.h>
    #include <stddef.h>
    #include "igmk.h"
    int main(void) {
        printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(igmk_system_t), sizeof(igmk_protocol_t), sizeof(igmk_result_t),
               offsetof(igmk_system_t, evfactor), offsetof(igmk_protocol_t, bond_prefactor),
               offsetof(igmk_result_t, md_seconds), offsetof(igmk_protocol_t, relax_mdsteps));
        return 0;
    }'''
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "t"), os.path.join(d, "t.c")])
        out = subprocess.check_output([os.path.join(d, "t")]).split()
    got = [int(v) for v in out]
    want = [C.sizeof(_abi.System), C.sizeof(_abi.Protocol), C.sizeof(_abi.Result), _abi.System.evfactor.offset,
            _abi.Protocol.bond_prefactor.offset, _abi.Result.md_seconds.offset, _abi.Protocol.relax_mdsteps.offset]
    assert got == want


def test_version_and_error_string_without_gpu():
    lib = _abi.load()
    assert lib.igmk_version() >= 100
    assert lib.igmk_device_count() >= 0
    # an argument error must not need a device
    assert lib.igmk_clear(None) != 0
    assert b"NULL" in lib.igmk_last_error()
