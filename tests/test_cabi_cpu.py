"""CPU-only: the C-ABI library builds for sm_100a, loads, exports every symbol
include/reth_replay.h declares, and fails loudly without a GPU (no fallback)."""
import ctypes as C
import re
import subprocess

import pytest

import __graft_entry__ as entry
from reth_b200 import _cabi


@pytest.fixture(scope="module", autouse=True)
def built():
    entry.build()


def _declared():
    text = open(_cabi.HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(reth_(?:replay|ipc|lz4f)_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound():
    names = _declared()
    assert len(names) >= 49 and "reth_ipc_open" in names and "reth_lz4f_decompress" in names
    syms = subprocess.check_output(["nm", "-D", "--defined-only", _cabi.LIB_PATH]).decode()
    L = _cabi.lib()
    for n in names:
        assert re.search(rf"\bT {n}\b", syms), f"{n} not exported"
        assert hasattr(L, n)
        assert n in _cabi.SIGNATURES or n in _cabi.LZ4F_SIGNATURES or n == "reth_replay_last_error", f"{n} has no ctypes signature"
    assert L.reth_replay_abi_version() == 3


def test_library_is_sm100a_with_lineinfo():
    out = subprocess.check_output(["cuobjdump", "-lelf", _cabi.LIB_PATH]).decode()
    assert "sm_100a" in out


def test_config_struct_matches_header():
    # 8 + 4(+4 pad) + 16*8 + 7*4 + 5*4 = 192
    assert C.sizeof(_cabi.Config) == 256      # 192 + out_cast[16]
    assert C.sizeof(_cabi.Stats) == 64


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from reth_b200 import ReplayError, ReplayShard
    with pytest.raises(ReplayError, match="no CUDA device"):
        ReplayShard(128, [], 8)


def test_product_does_not_import_oracle():
    import os
    root = os.path.dirname(_cabi.__file__)
    for dirpath, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "per_oracle" not in src and "from oracle" not in src and "import oracle" not in src, f
