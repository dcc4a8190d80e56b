#!/usr/bin/env python3
"""Cut small excerpts of the reference's DSD fixtures into tests/golden/fixtures.npz.

Run in the build container (needs /root/reference); the GPU box only sees the committed .npz.
Every entry keeps the container-order payload bytes of the first blocks of a fixture plus its
format, so the parity tests can run the same inputs through the oracle and the CUDA path.
Also records sha256 digests of the ORACLE's outputs on these excerpts (golden_digests.json) so a
silent change of the oracle is caught.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("D2P_REFERENCE", "/root/reference")

# name -> (path, payload offset, channels, layout, lsb_first, dsd_rate, blocks-per-channel kept)
FIX = {
    "impulse_mono_toggle": ("test/impulse_mono_toggle.dsd", 0, 1, "P", True, 1, 1),
    "1kHz_mono_p": ("test/1kHz_mono_p.dsf", 92, 1, "P", True, 1, 24),
    "1kHz_stereo_128": ("test/1kHz_stereo_128.dsf", 92, 2, "P", True, 2, 16),
    "pinknoise_stereo_128": ("test/pinknoise_stereo_128.dsf", 92, 2, "P", True, 2, 24),
    "pinknoise_mono_128": ("test/pinknoise_mono_128.dsf", 92, 1, "P", True, 2, 16),
    "1kHz_stereo_i": ("test/1kHz_stereo_i.dsd", 0, 2, "I", False, 1, 16),
}


def main():
    from oracle import orclib
    out, meta, digests = {}, {}, {}
    for name, (rel, off, ch, layout, lsb, rate, blocks) in FIX.items():
        raw = np.fromfile(os.path.join(REF, rel), dtype=np.uint8)[off:]
        n = min(raw.size, blocks * 4096 * ch)
        out[name] = raw[:n].copy()
        meta[name] = {"channels": ch, "layout": layout, "lsb_first": lsb, "dsd_rate": rate, "source": rel,
                      "payload_offset": off, "bytes": int(n)}
        for out_rate, bits in ((352800, 24), (176400, 16), (96000, 24), (192000, 32)):
            r = orclib.convert(out[name], dsd_rate=rate, out_rate=out_rate, bits=bits, dither="X", channels=ch,
                               layout=layout, lsb_first=lsb)
            digests[f"{name}/{out_rate}/{bits}"] = {"frames": r.frames[0],
                                                    "pcm_sha256": hashlib.sha256(r.pcm.tobytes()).hexdigest(),
                                                    "f64_sha256": hashlib.sha256(r.f64.tobytes()).hexdigest()}
    np.savez_compressed(os.path.join(HERE, "fixtures.npz"), **out)
    json.dump(meta, open(os.path.join(HERE, "fixtures_meta.json"), "w"), indent=1)
    json.dump(digests, open(os.path.join(HERE, "golden_digests.json"), "w"), indent=1)
    print({k: v.size for k, v in out.items()})


if __name__ == "__main__":
    main()
