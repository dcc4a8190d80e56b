#!/usr/bin/env python3
"""stream_push.py -- the reference's call pattern through the streaming context: one 4096-byte block per channel
per d2p_push (ConversionContext::process_blocks), d2p_pull after every block group.  Wall-clock DSD samples/s for a
few push thresholds (0 = ship every push: one copy + one launch per 8 KiB)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import ctypes as C
    import numpy as np
    import dsd2dxd_b200 as d2p
    from dsd2dxd_b200 import _capi
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    per_ch = int(seconds * 2822400 * 4 / 8) // 4096 * 4096
    raw = np.random.default_rng(1).integers(0, 256, size=2 * per_ch, dtype=np.uint8)
    cfg = d2p.make_config(dsd_rate=4, out_rate=352800, bits=24, dither="T", channels=2)
    L = _capi.lib()
    out = np.zeros(d2p.frames_for_bytes(cfg, per_ch) * 6 + 64, dtype=np.uint8)
    for thr in (0, 1 << 20, 8 << 20, 64 << 20):
        conv = d2p.Converter(cfg)
        conv.set_push_threshold(thr)
        base, got, at = raw.ctypes.data, C.c_uint64(), 0
        n = raw.size if thr else min(raw.size, 4096 * 2 * 2000)     # the unbatched mode is slow: bounded sample
        t0 = time.perf_counter()
        for off in range(0, n, 4096):
            _capi.check(L.d2p_push(conv._h, C.c_void_p(base + off), 4096))
            if off & 4096:                                           # after the last channel of a block group
                _capi.check(L.d2p_pull(conv._h, C.c_void_p(out.ctypes.data + at), out.size - at, C.byref(got)))
                at += got.value * 6
        conv.finish()
        _capi.check(L.d2p_pull(conv._h, C.c_void_p(out.ctypes.data + at), out.size - at, C.byref(got)))
        dt = time.perf_counter() - t0
        conv.close()
        print(json.dumps({"push_threshold": thr, "pushed_bytes": n, "seconds": dt, "dsd_samples_per_s": n * 8 / dt,
                          "realtime_factor": n * 8 / dt / (2 * 11289600.0)}), flush=True)


if __name__ == "__main__":
    main()
