"""CPU: the C-ABI shared library loads without a GPU and exports every symbol include/dualhyp.h declares;
the ctypes binding types each of them; compute entries fail loudly (no fallback) when there is no device."""
import ctypes
import os
import re

import pytest
import torch

from dualhyp_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "dualhyp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dh_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = header_symbols()
    assert len(names) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dualhyp.h but not exported by libdualhyp.so"
    assert sorted(_lib.SIGNATURES) == names, "ctypes binding and header disagree on the entry-point list"
    assert _lib.load().dh_abi_version() == 4


def test_gemm_args_struct_matches_header_layout():
    # 4 pointer/int64 pairs, 14 int32/float scalars, 3 pointer/int64 pairs, 6 pointers (ABI v2 added m_limit)
    assert ctypes.sizeof(_lib.GemmArgs) == 4 * 16 + 14 * 4 + 3 * 16 + 8 * 8          # ABI v4: + workspace, workspace_bytes


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    with pytest.raises(_lib.DualHypError):
        _lib.init(0)
    from dualhyp_b200 import ops
    with pytest.raises(_lib.DualHypError):
        ops.rmsnorm_fwd(torch.zeros(2, 256), torch.ones(256), torch.zeros(2, 256, dtype=torch.bfloat16), None, 1e-5)
