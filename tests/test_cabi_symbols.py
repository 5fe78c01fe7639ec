"""CPU: the C-ABI library loads and exports every symbol include/explainn_b200.h declares, and the
ctypes binding covers each of them (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "explainn_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(explainn_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_entry_points():
    names = declared_symbols()
    assert "explainn_conv_pool_fwd" in names and "explainn_last_error" in names
    assert len(names) >= 15


def test_library_exports_every_declared_symbol():
    from explainn_b200 import _cabi
    lib = ctypes.CDLL(_cabi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), "libexplainn_b200.so does not export %s" % name
        assert name in _cabi.SIGNATURES, "ctypes binding lacks %s" % name
    assert _cabi.lib.explainn_abi_version() == 1


def test_size_queries_and_error_string():
    from explainn_b200 import _cabi
    assert _cabi.lib.explainn_conv_tables_floats(300, 19) == 10 * 5 * 256 * 32
    assert _cabi.lib.explainn_conv_wq_floats(100, 19) == 4 * 4 * 19 * 32
    assert isinstance(_cabi.last_error(), str)


def test_no_cpu_path():
    import torch
    from explainn_b200 import networks
    m = networks.ExplaiNN(4, 40, 2, 8).eval()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(2, 4, 40))


def test_state_dict_layout_matches_reference():
    import torch
    from explainn_b200 import networks
    from oracle import torch_port as tp
    torch.manual_seed(0)
    a = networks.ExplaiNN(5, 60, 3, 19)
    torch.manual_seed(0)
    b = tp.PortExplaiNN(5, 60, 3, 19)
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa.keys()) == list(sb.keys()) and len(sa) == 23
    for k in sa:
        assert sa[k].shape == sb[k].shape and torch.equal(sa[k], sb[k]), k     # same default init stream
    assert a._options["num_cnns"] == 5 and a._options["filter_size"] == 19
    with pytest.raises(RuntimeError):
        a.load_state_dict(networks.ExplaiNN(6, 60, 3, 19).state_dict())
