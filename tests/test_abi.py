"""CPU-side checks of the drop-in boundary: the shared library loads, exports
every symbol include/d4est_b200.h declares, and refuses to compute without a
GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT
from disco4est_b200 import EXPORTS, LIB_PATH, MeshDesc, load_library


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "d4est_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(d4b_[a-z0-9_]+)\s*\(", text)) - {"d4b_abort_hook_t"})


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB_PATH), "run __graft_entry__.build() first"
    L = load_library()
    declared = _declared_symbols()
    assert declared, "header parse failed"
    for name in declared:
        assert hasattr(L, name), "libd4est_b200.so does not export %s" % name
    assert sorted(EXPORTS) == declared


def test_abi_version_and_desc_size():
    L = load_library()
    assert L.d4b_abi_version() == 1
    # field-for-field mirror of d4b_mesh_desc_t: 44 fields; pointer-sized alignment
    assert C.sizeof(MeshDesc) % 8 == 0


def test_no_cpu_fallback():
    L = load_library()
    if L.d4b_device_count() > 0:
        pytest.skip("GPU present")
    from disco4est_b200 import Mesh

    with pytest.raises(RuntimeError):
        Mesh({})
