#!/usr/bin/env python3
"""make_vectors.py -- run the REAL reference (clone206/dsd2dxd, built with cargo) over its own fixtures and write the
vectors that pin this repository's oracle: tests/golden/reference_vectors/{index.json, *.pcm, *.f64}.

The build container of this repository has no Rust toolchain and the reference's `dsd_reader` path dependency is missing
from the mount, so these vectors cannot be produced here: DESIGN.md says PARITY UNPINNED until someone
with `cargo` runs this once and commits the result.  Steps (see README.md in this directory):

    git clone https://github.com/clone206/dsd2dxd && cd dsd2dxd
    patch -p1 < /path/to/dsd2pcm-b200/tools/ref_dump/dump_f64.patch
    cargo build --release
    python /path/to/dsd2pcm-b200/tools/ref_dump/make_vectors.py --ref . --bin target/release/dsd2dxd

Every case runs the reference CLI with dither off (-d X: the only reproducible setting) and stdout output
(-o S), so the PCM bytes are pack_int / pack_float's, and with D2P_DUMP_F64 set, so the pre-dither f64 samples and the
reader's block sizes come out as well.  tests/test_reference_vectors.py consumes the result (oracle on CPU, CUDA path on
the GPU box): bit-exact PCM, bit-exact f64, equal frame counts.
"""
import argparse
import hashlib
import json
import os
import shutil
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

# fixture -> CLI flags that describe it (containers describe themselves) and the format the tests need
FIXTURES = {
    "test/1kHz_mono_p.dsf": ([], dict(container=True)),
    "test/1kHz_mono_128.dsf": ([], dict(container=True)),
    "test/1kHz_stereo_128.dsf": ([], dict(container=True)),
    "test/pinknoise_mono_128.dsf": ([], dict(container=True)),
    "test/pinknoise_stereo_128.dsf": ([], dict(container=True)),
    "id3_test/1kHz_mono_brokenid3.dsf": ([], dict(container=True)),
    "id3_test/dff/1kHz_stereo_i.dff": ([], dict(container=True)),
    "id3_test/dff/1kHz_stereo_i_brokenid3.dff": ([], dict(container=True)),
    "test/1kHz_mono_p.dsd": (["-c", "1", "-f", "P", "-e", "L", "-i", "1"], dict(channels=1, layout="P", lsb_first=True, dsd_rate=1)),
    "test/1kHz_stereo_i.dsd": (["-c", "2", "-f", "I", "-e", "M", "-i", "1"], dict(channels=2, layout="I", lsb_first=False, dsd_rate=1)),
    "test/impulse_mono_toggle.dsd": (["-c", "1", "-f", "P", "-e", "L", "-i", "1"], dict(channels=1, layout="P", lsb_first=True, dsd_rate=1)),
}
# (out_rate, filter) per DSD rate: SURVEY A2 (integer path) + A6 (L/M path)
COMBOS = {
    1: [(352800, "D"), (352800, "X"), (352800, "E"), (176400, "X"), (176400, "E"), (88200, "X"), (88200, "E"), (44100, "E"),
        (96000, "E"), (192000, "E"), (384000, "E")],
    2: [(705600, "E"), (352800, "E"), (352800, "C"), (352800, "X"), (176400, "E"), (176400, "C"), (88200, "E"), (88200, "C"),
        (96000, "E"), (192000, "E"), (384000, "E")],
}
BITS = [16, 20, 24, 32]


def dsd_rate_of(path):
    """DSF: fmt.sample_rate at byte 56 (LE); DSDIFF: PROP/FS chunk (BE).  Raw files carry it in FIXTURES."""
    b = open(path, "rb").read(1 << 16)
    if b[:4] == b"DSD ":
        return int.from_bytes(b[56:60], "little") // 2822400
    at = b.index(b"FS  ")
    return int.from_bytes(b[at + 12:at + 16], "big") // 2822400


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", required=True, help="checkout of clone206/dsd2dxd (holds test/ and id3_test/)")
    ap.add_argument("--bin", required=True, help="the built reference CLI (target/release/dsd2dxd), patched with dump_f64.patch")
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "reference_vectors"))
    ap.add_argument("--bits", type=int, nargs="*", default=BITS)
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    cases = []
    for rel, (flags, fmt) in FIXTURES.items():
        src = os.path.join(args.ref, rel)
        rate = fmt.get("dsd_rate") or dsd_rate_of(src)
        for out_rate, filt in COMBOS[rate]:
            for bits in args.bits:
                tmp = tempfile.mkdtemp()
                env = dict(os.environ, D2P_DUMP_F64=tmp)
                cmd = [args.bin] + flags + ["-b", str(bits), "-t", filt, "-d", "X", "-r", str(out_rate), "-o", "S", "-q"]
                if fmt.get("container"):
                    r = subprocess.run(cmd + [src], env=env, capture_output=True)
                else:
                    r = subprocess.run(cmd + ["-"], env=env, stdin=open(src, "rb"), capture_output=True)
                tag = hashlib.sha1(f"{rel}/{out_rate}/{filt}/{bits}".encode()).hexdigest()[:16]
                case = {"fixture": rel, "format": fmt, "dsd_rate": rate, "out_rate": out_rate, "filter": filt, "bits": bits,
                        "dither": "X", "returncode": r.returncode, "stderr": r.stderr.decode(errors="replace")[-400:]}
                if r.returncode == 0:
                    open(os.path.join(args.out, tag + ".pcm"), "wb").write(r.stdout)
                    chans = sorted(f for f in os.listdir(tmp) if f.endswith(".f64"))
                    if bits == 24:                                # the f64 samples do not depend on the bit depth: keep one copy
                        with open(os.path.join(args.out, tag + ".f64"), "wb") as out:
                            for c in chans:                       # channel-major: ch0 frames, ch1 frames, ...
                                out.write(open(os.path.join(tmp, c), "rb").read())
                        case["f64"] = tag + ".f64"
                    case.update(pcm=tag + ".pcm", channels=len(chans),
                                frames=os.path.getsize(os.path.join(tmp, chans[0])) // 8 if chans else 0,
                                reader_blocks=open(os.path.join(tmp, "reader_blocks.txt")).read().split("\n")[:4]
                                + ["..."] + open(os.path.join(tmp, "reader_blocks.txt")).read().strip().split("\n")[-2:])
                cases.append(case)
                shutil.rmtree(tmp)
    json.dump({"made_by": "tools/ref_dump/make_vectors.py", "reference": "clone206/dsd2dxd + dump_f64.patch", "cases": cases},
              open(os.path.join(args.out, "index.json"), "w"), indent=1)
    print(f"{len(cases)} cases -> {args.out}")


if __name__ == "__main__":
    main()
