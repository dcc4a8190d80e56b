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
    assert L.d2p_abi_version() == 2
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


# ---- the reference-side binding (integration/dsd2pcm-b200-sys) stays in step with the headers ----------------
def _headers():
    text = "".join(open(os.path.join(ROOT, "include", h)).read() for h in ("dsd2pcm_b200.h", "dsd2pcm_b200_files.h"))
    return re.sub(r"/\*.*?\*/", "", text, flags=re.S)


def _rust():
    return open(os.path.join(ROOT, "integration", "dsd2pcm-b200-sys", "src", "lib.rs")).read()


def test_rust_ffi_crate_declares_exactly_the_header_functions():
    rs = _rust()
    assert sorted(set(re.findall(r"pub fn (d2p_[a-z0-9_]+)\s*\(", rs))) == declared_symbols()


C_TO_RUST = {"int32_t": "i32", "uint32_t": "u32", "int64_t": "i64", "uint64_t": "u64", "double": "f64", "float": "f32",
             "const void*": "*const c_void", "void*": "*mut c_void", "const char*": "*const c_char", "double*": "*mut f64", "d2p_container_info": "d2p_container_info"}


def _c_struct_fields(name):
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), _headers(), flags=re.S).group(1)
    fields = []
    for decl in body.split(";"):
        decl = " ".join(decl.split())
        if not decl:
            continue
        m = re.match(r"(const void\*|const char\*|void\*|double\*|[a-z0-9_]+)\s+(.*)", decl)
        ctype, names = m.group(1), m.group(2)
        for nm in names.split(","):
            nm = nm.strip()
            arr = re.findall(r"\[(\d+)\]", nm)
            fields.append((re.sub(r"\[.*", "", nm), ctype, [int(a) for a in arr]))
    return fields


def _rust_struct_fields(name):
    body = re.search(r"pub struct %s \{(.*?)\n\}" % name, _rust(), flags=re.S).group(1)
    return [(m.group(1), m.group(2).strip()) for m in re.finditer(r"pub ((?:r#)?\w+): ([^,\n]+),", body)]


@pytest.mark.parametrize("name", ["d2p_config", "d2p_plan_info", "d2p_job", "d2p_block_job", "d2p_container_info",
                                  "d2p_file_options", "d2p_file_result", "d2p_flac_picture"])
def test_rust_structs_mirror_the_c_layout(name):
    cf, rf = _c_struct_fields(name), _rust_struct_fields(name)
    assert len(cf) == len(rf), (cf, rf)
    for (cn, ct, arr), (rn, rt) in zip(cf, rf):
        assert rn == cn or (cn, rn) in (("in", "input"), ("type", "r#type"))      # Rust keywords
        if ct == "char":
            want = "c_char"
        else:
            want = C_TO_RUST[ct]
        for a in reversed(arr):
            want = f"[{want}; {a}]"
        assert rt == want, (name, cn, rt, want)


def test_rust_constants_match_the_header_enums():
    hs, rs = _headers(), _rust()
    c_vals = {}
    for m in re.finditer(r"\b(D2P_[A-Z0-9_]+)\s*=\s*('.'|\d+)", hs):
        v = m.group(2)
        c_vals[m.group(1)] = ord(v[1]) if v.startswith("'") else int(v)
    r_vals = {}
    for m in re.finditer(r"pub const (D2P_[A-Z0-9_]+): \w+ = (b'.' as i32|\d+);", rs):
        v = m.group(2)
        r_vals[m.group(1)] = ord(v[2]) if v.startswith("b'") else int(v)
    r_vals.pop("D2P_ABI_VERSION")
    assert r_vals == c_vals


def test_ctypes_structs_have_the_c_field_order():
    for cname, cls in (("d2p_config", _capi.Config), ("d2p_container_info", _capi.ContainerInfo),
                       ("d2p_file_options", _capi.FileOptions)):
        assert [f[0] for f in cls._fields_] == [c[0] for c in _c_struct_fields(cname)], cname
