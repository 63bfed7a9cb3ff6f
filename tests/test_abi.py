"""CPU: the C-ABI library loads and exports every symbol include/b200ir.h declares (no compute)."""
import ctypes
import os
import re

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "b200ir.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ir_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    import b200ir
    names = _declared()
    assert len(names) >= 10
    lib = ctypes.CDLL(b200ir._lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200ir.h but not exported"
        assert n in b200ir._lib.SIGNATURES, f"{n} has no ctypes signature in _lib.py"
    assert set(b200ir._lib.SIGNATURES) <= set(names), "ctypes binds symbols the header does not declare"


def test_version_and_error_string():
    import b200ir
    assert b200ir._lib.lib.b200ir_version() >= 100
    # bad-argument call (NULL pointers) must be rejected on the host, before any CUDA call
    rc = b200ir._lib.lib.b200ir_dwconv3x3_fwd(None, 0, None, None, None, 0, 0, 1, 1, 1, 1, None)
    assert rc == -1
    assert b"dwconv3x3" in b200ir._lib.lib.b200ir_last_error_string()


def test_modules_refuse_cpu_tensors():
    import pytest
    import torch
    import b200ir
    blk = b200ir.TransformerBlock(48, 1, 2.66, False, "WithBias")
    with pytest.raises(ValueError, match="CUDA"):
        with torch.no_grad():
            blk(torch.zeros(1, 48, 8, 8))


def test_state_dict_contract():
    import b200ir
    from oracle import restormer_oracle as O
    net = b200ir.Restormer()
    sd = O.synth_state_dict(seed=0)
    assert list(sd.keys()) == list(net.state_dict().keys())
    net.load_state_dict(sd, strict=True)
    blk = b200ir.TransformerBlock(96, 2, 2.66, True, "BiasFree")
    assert list(blk.state_dict().keys()) == list(O.synth_block_params(96, 2, 2.66, True, "BiasFree", 1).keys())
