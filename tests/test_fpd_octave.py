"""The one libm-dependent expression of the hot path: Dither::fpdither's octave `s.abs().log2().floor() as i32`
(rd/dither.rs:102-104).  The oracle evaluates it with the host's glibc log2 (which is what a Linux build of the Rust
reference links).  libdevice's log2 is NOT usable for it: it is inexact at powers of two and lands in the wrong octave on
and next to every octave edge (measured: 206 of 2600 edge samples), so the kernels derive the octave from the exponent
field plus the round-up rule of a correctly rounded log2 (d2p_kernels.cu: fpd_exponent).  This test walks every octave
edge from 2^-70 to 2^39: the power of two, the 800 doubles below it (the round-up zone is at most 709 wide) and the 8
above, both signs, plus zero / subnormal / inf / NaN."""
import ctypes as C

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import _capi
from oracle import orclib


def boundary_samples():
    vals = []
    for k in range(-70, 40):
        p = np.ldexp(1.0, k)
        below = p - np.arange(0, 801, dtype=np.float64) * np.spacing(np.nextafter(p, 0.0))   # 2^k and 800 doubles below it
        above = p + np.arange(1, 9, dtype=np.float64) * np.spacing(p)
        vals += [below, above]
    vals.append(np.array([0.0, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, np.inf, np.nan, 0.75, 1.5, 3.0]))
    v = np.concatenate(vals)
    return np.concatenate([v, -v])


def oracle_exponents(s):
    L = orclib.lib()
    L.orc_fpd_exponent.argtypes = [C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_int32)]
    L.orc_fpd_exponent.restype = None
    e = np.zeros(s.size, dtype=np.int32)
    L.orc_fpd_exponent(s.ctypes.data_as(C.POINTER(C.c_double)), s.size, e.ctypes.data_as(C.POINTER(C.c_int32)))
    return e


def test_oracle_octave_is_the_exponent_field_away_from_the_boundaries():
    s = np.array([0.75, 1.0, 1.5, 2.0, 3.0, 0.5, 2.0 ** -20 * 1.25, 1e-300], dtype=np.float64)
    assert list(oracle_exponents(s)) == [-1, 0, 0, 1, 1, -1, -20, -997]
    assert list(oracle_exponents(np.array([0.0, np.nan, np.inf]))) == [-2147483648, 0, 2147483647]


@pytest.mark.gpu
def test_device_octave_equals_host_libm_at_every_power_of_two_boundary():
    s = boundary_samples()
    want = oracle_exponents(s)
    got = np.zeros(s.size, dtype=np.int32)
    _capi.check(_capi.lib().d2p_debug_fpd_exponent(s.ctypes.data_as(C.POINTER(C.c_double)), s.size,
                                                   got.ctypes.data_as(C.POINTER(C.c_int32)), 0))
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, [(float.hex(float(s[i])), int(got[i]), int(want[i])) for i in bad[:10]]
