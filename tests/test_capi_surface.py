"""The C-ABI library loads on a CPU-only box and exports every symbol include/*.h declare.
Host-only entries (planning, tap tables, error strings) are exercised; compute entries must fail
loudly with D2P_ERR_NO_DEVICE when no GPU is present (there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = "".join(open(os.path.join(ROOT, "include", h)).read() for h in ("dsd2pcm_b200.h", "dsd2pcm_b200_files.h"))
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(d2p_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    names = declared_symbols()
    assert len(names) >= 26
    L = C.CDLL(_capi.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(_capi.EXPORTS) == names          # the ctypes mirror binds exactly the header's surface


def test_abi_version_and_error_strings():
    L = _capi.lib()
    assert L.d2p_abi_version() == 1
    assert b"Taps not found" in L.d2p_strerror(2)
    assert b"Unsupported L/M" in L.d2p_strerror(3)
    assert b"Unsupported bit depth" in L.d2p_strerror(4)


def test_plan_is_host_only_and_mirrors_reference_setup():
    info = d2p.plan(d2p.make_config(dsd_rate=4, out_rate=352800))
    assert info["tables"] == ["HTAPS_DSD256_32TO1_EQ"] and info["half_taps"] == [151] and info["frames_cap"] == 1024
    info = d2p.plan(d2p.make_config(dsd_rate=4, out_rate=192000, bits=32))
    assert info["tables"] == ["HTAPS_DDRX10_21TO1_EQ", "HTAPS_2_68MHZ_14TO1_EQ"]
    assert info["decims"] == [21, 14] and info["upsample_ratio"] == 5 and info["scale_factor"] == 5.0
    assert info["frames_cap"] == 574
    with pytest.raises(d2p.D2PError) as e:                      # F11: X and D have no 64:1 taps
        d2p.plan(d2p.make_config(dsd_rate=2, out_rate=88200, filt="X"))
    assert "Taps not found for decimation ratio 64" in str(e.value)
    with pytest.raises(d2p.D2PError):
        d2p.plan(d2p.make_config(bits=12))


def test_tap_tables_are_served_by_the_product_library():
    names = d2p.tap_table_names()
    assert len(names) == 33
    t = d2p.tap_table("HTAPS_DSD64_8TO1_EQ")
    assert t.size == 54 and abs(2 * t.sum() - 1.0) < 0.01


def test_frames_for_bytes_closed_form():
    cfg = d2p.make_config(dsd_rate=1, out_rate=352800)
    assert d2p.frames_for_bytes(cfg, 1764000) == 1763947
    cfg = d2p.make_config(dsd_rate=4, out_rate=352800)
    assert d2p.frames_for_bytes(cfg, 846720000) == 211679850      # cfg5


@pytest.mark.skipif(d2p.device_count() > 0, reason="checks the no-GPU failure mode")
def test_compute_entries_fail_loudly_without_a_gpu():
    with pytest.raises(d2p.D2PError) as e:
        d2p.Converter(d2p.make_config())
    assert e.value.status == 5 and "no CPU path" in str(e.value)
    with pytest.raises(d2p.D2PError):
        d2p.convert_host(d2p.make_config(), np.zeros(8192, dtype=np.uint8))
