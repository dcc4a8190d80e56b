#!/usr/bin/env python3
"""run_path.py -- run one conversion configuration on device-resident synthetic DSD and time it.

A profiling / exploration helper (bench.py is the contract benchmark).  Every BASELINE.json config can be
reproduced in shape, scaled by --seconds:
    python tools/run_path.py --dsd-rate 4 --out-rate 192000 --bits 32 --seconds 60
Prints one JSON line: kernel ms (CUDA events, median of --reps), DSD samples/s, LUT look-ups/s.
Use under `ncu -k regex:lutfir|lm_kernel` to capture the kernel of any configuration.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dsd-rate", type=int, default=4)
    ap.add_argument("--out-rate", type=int, default=352800)
    ap.add_argument("--filter", default="E")
    ap.add_argument("--bits", type=int, default=24)
    ap.add_argument("--dither", default="X")
    ap.add_argument("--channels", type=int, default=2)
    ap.add_argument("--layout", default="P")
    ap.add_argument("--msb-first", action="store_true")
    ap.add_argument("--seconds", type=float, default=60.0, help="audio seconds per file")
    ap.add_argument("--files", type=int, default=1)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()

    import torch
    import dsd2dxd_b200 as d2p

    dev = torch.device("cuda", 0)
    per_ch = int(args.seconds * 2822400 * args.dsd_rate / 8) // 4096 * 4096
    in_bytes = per_ch * args.channels
    cfg = d2p.make_config(dsd_rate=args.dsd_rate, out_rate=args.out_rate, filt=args.filter, bits=args.bits,
                          dither=args.dither, channels=args.channels, layout=args.layout,
                          lsb_first=not args.msb_first, block_size=4096, dither_seed=1)
    info = d2p.plan(cfg)
    frames = d2p.frames_for_bytes(cfg, per_ch)
    out_bytes = frames * args.channels * info["bytes_per_sample"]
    batch = d2p.Batch(cfg)
    gen = torch.Generator(device=dev)
    gen.manual_seed(0x5D2DDD)
    d_in = [torch.randint(0, 256, (in_bytes,), dtype=torch.uint8, device=dev, generator=gen) for _ in range(args.files)]
    d_out = [torch.empty(out_bytes + 16, dtype=torch.uint8, device=dev) for _ in range(args.files)]
    jobs = batch.make_jobs(dict(d_in=i.data_ptr(), in_bytes=in_bytes, d_out=o.data_ptr(), out_cap_bytes=o.numel(),
                                dither_seed=k) for k, (i, o) in enumerate(zip(d_in, d_out)))
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(args.warmup):
        batch.run(jobs, stream)
    torch.cuda.synchronize()
    times = []
    for _ in range(args.reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        batch.run(jobs, stream)
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    times.sort()
    ms = times[len(times) // 2]
    samples = args.files * in_bytes * 8
    print(json.dumps({"plan": info, "frames_per_file": frames, "kernel_ms": ms, "all_ms": times,
                      "dsd_samples_per_s": samples / (ms / 1e3),
                      "realtime_factor": args.files * args.seconds / (ms / 1e3),
                      "hbm_gbs_algorithmic": args.files * (in_bytes + out_bytes) / (ms / 1e3) / 1e9}))


if __name__ == "__main__":
    main()
