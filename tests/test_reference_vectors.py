"""Consumer of reference-run vectors (tests/golden/reference_vectors/, made by tools/ref_dump/make_vectors.py from the
REAL clone206/dsd2dxd built with cargo).  When the directory is absent -- it is, until someone with a Rust toolchain
runs the recipe in tools/ref_dump/README.md -- every test here skips with "PARITY UNPINNED": the oracle is then checked
by known answers only (tests/test_oracle_kat.py) and DESIGN.md says so.

With vectors present, per case (fixture x output rate x filter x bit depth, dither X):
  CPU : the oracle reproduces the reference's frame count, pre-dither f64 samples and PCM bytes, bit for bit;
  GPU : so does the CUDA path through the C ABI.
The DSF reader question of SURVEY 8c (stop at fmt.sample_count or pass whole blocks) is decided by trying both; the
winner must be the same for every DSF case and is reported.
"""
import json
import os

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from oracle import orclib

HERE = os.path.dirname(os.path.abspath(__file__))
VEC = os.path.join(HERE, "golden", "reference_vectors")
UNPINNED = "PARITY UNPINNED: no reference-run vectors in tests/golden/reference_vectors (recipe: tools/ref_dump/README.md)"


def cases():
    idx = os.path.join(VEC, "index.json")
    if not os.path.exists(idx):
        return None
    return [c for c in json.load(open(idx))["cases"] if c.get("returncode") == 0]


def payload_variants(ref_files, case):
    """-> [(label, payload bytes, format kwargs)]: the reader-contract hypotheses to try for this fixture."""
    path = ref_files[case["fixture"]]
    fmt = dict(case["format"])
    if not fmt.pop("container", False):
        return [("raw", np.fromfile(path, dtype=np.uint8), fmt)]
    info = d2p.container_open(path)
    data = np.fromfile(path, dtype=np.uint8)
    pay = data[info["data_offset"]: info["data_offset"] + info["data_bytes"]]
    layout = chr(info["in_layout"])
    kw = dict(channels=info["channels"], layout=layout, lsb_first=bool(info["lsb_first"]), dsd_rate=info["dsd_rate"])
    if layout != "P":
        return [("dff", pay, kw)]
    kw["block_size"] = info["block_size"]
    out = [("dsf-whole-blocks", pay, kw)]
    per_ch = pay.size // info["channels"]
    limit = min((info["sample_count"] + 7) // 8, per_ch)
    if limit < per_ch:                                            # stop at sample_count: ragged last group, runs back to back
        bs, C = info["block_size"], info["channels"]
        full, tail = limit // bs, limit % bs
        last = pay[full * bs * C: (full + 1) * bs * C]
        cut = np.concatenate([pay[: full * bs * C]] + [last[c * bs: c * bs + tail] for c in range(C)])
        out.append(("dsf-stop-at-sample-count", cut, kw))
    return out


def load_case(case):
    pcm = np.fromfile(os.path.join(VEC, case["pcm"]), dtype=np.uint8)
    f64 = None
    if "f64" in case:
        f64 = np.fromfile(os.path.join(VEC, case["f64"]), dtype="<f8").reshape(case["channels"], -1)
    return pcm, f64


def check(case, ref_files, convert):
    pcm, f64 = load_case(case)
    winners = []
    for label, pay, fmt in payload_variants(ref_files, case):
        got_pcm, got_f64, frames = convert(pay, fmt, case, f64 is not None)
        if frames == case["frames"] and np.array_equal(got_pcm, pcm) and (f64 is None or np.array_equal(got_f64, f64)):
            winners.append(label)
    assert winners, (case["fixture"], case["out_rate"], case["filter"], case["bits"], "differs from the reference")
    return winners[0]


def test_oracle_reproduces_the_reference(ref_files):
    cs = cases()
    if cs is None:
        pytest.skip(UNPINNED)

    def convert(pay, fmt, case, want_f64):
        r = orclib.convert(pay, want_f64=want_f64, out_rate=case["out_rate"], filt=case["filter"], bits=case["bits"], dither="X", **fmt)
        return r.pcm, r.f64, r.frames[0]

    contract = {check(c, ref_files, convert) for c in cs}
    dsf = {w for w in contract if w.startswith("dsf-")}
    assert len(dsf) <= 1, f"the DSF reader contract is not the same for every case: {dsf}"
    print("reader contract the reference follows:", sorted(contract))


@pytest.mark.gpu
def test_cuda_path_reproduces_the_reference(ref_files):
    cs = cases()
    if cs is None:
        pytest.skip(UNPINNED)

    def convert(pay, fmt, case, want_f64):
        cfg = d2p.make_config(out_rate=case["out_rate"], filt=case["filter"], bits=case["bits"], dither="X", keep_f64=want_f64, **fmt)
        conv = d2p.Converter(cfg)
        conv.push(pay)
        conv.finish()
        f64 = np.stack([conv.peek_f64(c) for c in range(cfg.channels)]) if want_f64 else None
        pcm = conv.pull()
        conv.close()
        return pcm, f64, pcm.size // (cfg.channels * d2p.plan(cfg)["bytes_per_sample"])

    for c in cs:
        check(c, ref_files, convert)


def test_unpinned_state_is_said_out_loud():
    """DESIGN.md must carry the words PARITY UNPINNED exactly as long as no reference-run vectors are committed."""
    design = open(os.path.join(os.path.dirname(HERE), "DESIGN.md")).read()
    assert ("PARITY UNPINNED" in design) == (cases() is None)


def test_consumer_accepts_vectors_in_the_dump_layout(ref_files, tmp_path, monkeypatch):
    """The consumer itself, exercised on vectors in the exact layout tools/ref_dump/make_vectors.py writes -- produced here
    by the oracle (so this proves the plumbing, not parity): a pass, and a detected one-sample difference."""
    import sys
    mod = sys.modules[__name__]
    out = tmp_path / "reference_vectors"
    out.mkdir()
    cases_ = []
    for rel, fmt, out_rate, bits in (("test/1kHz_mono_p.dsd", dict(channels=1, layout="P", lsb_first=True, dsd_rate=1), 352800, 24),
                                     ("id3_test/dff/1kHz_stereo_i.dff", dict(container=True), 96000, 32)):
        case = {"fixture": rel, "format": fmt, "out_rate": out_rate, "filter": "E", "bits": bits, "dither": "X", "returncode": 0}
        label, pay, kw = payload_variants(ref_files, case)[0]
        r = orclib.convert(pay, out_rate=out_rate, filt="E", bits=bits, dither="X", **kw)
        tag = f"case{len(cases_)}"
        r.pcm.tofile(out / (tag + ".pcm"))
        r.f64.astype("<f8").tofile(out / (tag + ".f64"))
        case.update(pcm=tag + ".pcm", f64=tag + ".f64", channels=r.f64.shape[0], frames=r.frames[0])
        cases_.append(case)
    json.dump({"cases": cases_}, open(out / "index.json", "w"))
    monkeypatch.setattr(mod, "VEC", str(out))

    def convert(pay, fmt, case, want_f64):
        r = orclib.convert(pay, want_f64=want_f64, out_rate=case["out_rate"], filt=case["filter"], bits=case["bits"], dither="X", **fmt)
        return r.pcm, r.f64, r.frames[0]

    assert [check(c, ref_files, convert) for c in cases()] == ["raw", "dff"]
    pcm = np.fromfile(out / "case0.pcm", dtype=np.uint8)
    pcm[1000] ^= 1
    pcm.tofile(out / "case0.pcm")
    with pytest.raises(AssertionError, match="differs from the reference"):
        check(cases()[0], ref_files, convert)
