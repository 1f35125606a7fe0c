"""CPU-side checks of the boundary: libbdq.so loads without a GPU and exports every symbol include/bdq.h declares;
argument validation that needs no device; host-side helpers."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "bdq.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bdq_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from bitdistiller_b200 import _lib
    lib = _lib.lib()
    names = _declared()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/bdq.h but not exported by libbdq.so"
    assert lib.bdq_version() >= 100


def test_zeros_width_matches_reference_rule():
    """calculate_zeros_width, qmodule.py:16-28 (values from SURVEY App. C)."""
    from bitdistiller_b200 import _lib
    from bitdistiller_b200.qmodule import calculate_zeros_width
    lib = _lib.lib()
    lib.bdq_zeros_width.argtypes = [ctypes.c_int64, ctypes.c_int, ctypes.c_int]
    for ic, g, wbit, want in [(4096, 128, 2, 2), (11008, 128, 2, 6), (12288, 128, 2, 6), (4096, 64, 2, 4),
                              (4096, 128, 4, 4), (4096, 32, 4, 16), (4096, 4096, 2, 1)]:
        assert lib.bdq_zeros_width(ic, g, wbit) == want
        assert calculate_zeros_width(ic, g, 32 // wbit) == want
    assert lib.bdq_zeros_width(4096, 48, 2) == -1
    with pytest.raises(NotImplementedError):
        calculate_zeros_width(4096, 48, 16)


def test_argument_validation_without_device():
    from bitdistiller_b200 import _lib
    lib = _lib.lib()
    i64 = ctypes.c_int64
    assert lib.bdq_int_fakequant_fwd(None, None, None, None, None, None, None, i64(100), 128, 2, 1, 0, 0, None) == -1
    assert b"multiple of group" in lib.bdq_last_error()
    assert lib.bdq_int_fakequant_fwd(None, None, None, None, None, None, None, i64(128), 128, 9, 1, 0, 0, None) == -5
    assert lib.bdq_int_fakequant_fwd(None, None, None, None, None, None, None, i64(0), 128, 2, 1, 0, 0, None) == 0
    assert lib.bdq_pack_wq(None, None, None, None, None, None, i64(4), i64(128), 128, 3, 1, None) == -5
    assert lib.bdq_gemv_w2(None, None, None, None, None, i64(1), i64(100), i64(16), 128, 0, None) == -1


def test_python_api_surface_matches_reference_names():
    import inspect

    from bitdistiller_b200 import qlinear, qmodule, quantizer
    sig = inspect.signature(quantizer.pseudo_quantize_tensor)
    assert list(sig.parameters) == ["w", "n_bit", "zero_point", "q_group_size", "inplace", "get_scale_zp"]
    assert [p.default for p in sig.parameters.values()][1:] == [8, True, -1, False, False]
    assert quantizer.SteInt2AsymQuantizer().q_group_size == 64 and quantizer.SteInt3AsymQuantizer().q_group_size == 128
    assert quantizer.SteN2F3Quantizer().q_group_size == 128
    sig = inspect.signature(qlinear.convertModelToQuant)
    assert sig.parameters["quant_type"].default == "clsq-n2f3" and sig.parameters["q_group_size"].default == 128
    with pytest.raises(ValueError):
        qlinear.QLinear(8, 8, quant_type="clsq-n2f3", device="meta")
    q = qlinear.QLinear(128, 64, bias=False, quant_type="int2-asym", q_group_size=64, device="meta")
    assert isinstance(q.weight_quantizer, quantizer.SteInt2AsymQuantizer)
    sig = inspect.signature(qmodule.WQLinear.__init__)
    assert list(sig.parameters)[1:] == ["w_bit", "group_size", "in_features", "out_features", "bias", "dev"]
    m = qmodule.WQLinear(2, 128, 4096, 4096, True, "meta")
    assert m.qweight.shape == (4096, 256) and m.qzeros.shape == (4096, 2) and m.scales.shape == (4096, 32)
    assert m.split_k_iters == 16 and m.offset == 0x03
