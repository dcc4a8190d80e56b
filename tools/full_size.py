#!/usr/bin/env python3
"""full_size.py -- BASELINE.json configs 3a, 3b and 4 at their FULL sizes on one B200, device-resident.

For each configuration: generate the 60-minute synthetic payload on the device, convert it in ONE launch
(CUDA-event time), convert it again as N time blocks that each see only their own input slice
(d2p_block_range + d2p_batch_run_blocks) and check the size-independent property
    digest(one pass) == digest(blocks concatenated)
with a 64-bit checksum of checksums computed on the device (no 20-50 GB copy back).  Also checks the frame count
against d2p_frames_for_bytes.  One JSON line per configuration.
    python tools/full_size.py [--minutes 60] [--blocks 8] [--only 3a,3b,4]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = {
    "3a": dict(dsd_rate=4, out_rate=176400, bits=20, channels=2, layout="P", lsb_first=True, dither="X",
               what="60 min DSD256 stereo -> 20-bit 176.4 kHz (HTAPS_DSD256_64TO1_EQ)"),
    "3b": dict(dsd_rate=4, out_rate=192000, bits=32, channels=2, layout="P", lsb_first=True, dither="X",
               what="60 min DSD256 stereo -> 32-bit float 192 kHz (L/M 5/294, two stages)"),
    "4": dict(dsd_rate=8, out_rate=352800, bits=24, channels=6, layout="I", lsb_first=False, dither="X",
              what="60 min DSD512 6-ch byte-interleaved MSB-first -> 24-bit 352.8 kHz (HTAPS_DSD512_64TO1_EQ)"),
    "4q": dict(dsd_rate=8, out_rate=384000, bits=24, channels=6, layout="I", lsb_first=False, dither="X",
               what="60 min DSD512 6-ch interleaved MSB-first -> '384 kHz' as written (L/M quirk path, SURVEY F8)"),
}


def digest(t, torch):
    """Order-sensitive 64-bit checksum of a uint8 device tensor: sum_i (i+1)*word_i over int64 words of 8 bytes, in
    1 GiB pieces (checksum of checksums, wrap-around arithmetic)."""
    n = t.numel() // 8 * 8
    total = 0
    piece = 1 << 30
    for off in range(0, n, piece):
        w = t[off:min(n, off + piece)].view(torch.int64)
        idx = torch.arange(1, w.numel() + 1, device=t.device, dtype=torch.int64)
        total = (total * 1000003 + int((w * idx).sum().item())) & 0xFFFFFFFFFFFFFFFF
    tail = t[n:].cpu().numpy().tobytes()
    return f"{total:016x}+{tail.hex()}"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--minutes", type=float, default=60.0)
    ap.add_argument("--blocks", type=int, default=8)
    ap.add_argument("--only", default="3a,3b,4")
    ap.add_argument("--shard", action="store_true",
                    help="under torchrun: every rank converts only ITS time blocks of the payload (own GPU, own slice); "
                         "rank 0 also converts the whole payload in one pass and compares the per-block digests")
    args = ap.parse_args()
    if args.shard:
        return shard_main(args)

    import torch
    import dsd2dxd_b200 as d2p
    from dsd2dxd_b200 import shard

    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream().cuda_stream
    for name in args.only.split(","):
        kw = dict(CONFIGS[name])
        what = kw.pop("what")
        C = kw["channels"]
        per_ch = int(args.minutes * 60 * 2822400 * kw["dsd_rate"] / 8) // 4096 * 4096
        in_bytes = per_ch * C
        cfg = d2p.make_config(block_size=4096, dither_seed=7, filt="E", **kw)
        info = d2p.plan(cfg)
        bps = info["bytes_per_sample"]
        frames = d2p.frames_for_bytes(cfg, per_ch)
        out_bytes = frames * C * bps
        gen = torch.Generator(device=dev)
        gen.manual_seed(0x5D2DDD)
        d_in = torch.empty(in_bytes, dtype=torch.uint8, device=dev)
        for off in range(0, in_bytes, 1 << 30):                  # 1 GiB pieces keep randint's temporaries small
            n = min(1 << 30, in_bytes - off)
            d_in[off:off + n] = torch.randint(0, 256, (n,), dtype=torch.uint8, device=dev, generator=gen)
        d_out = torch.zeros(out_bytes + 16, dtype=torch.uint8, device=dev)
        batch = d2p.Batch(cfg)
        jobs = batch.make_jobs([dict(d_in=d_in.data_ptr(), in_bytes=in_bytes, d_out=d_out.data_ptr(),
                                     out_cap_bytes=d_out.numel(), dither_seed=7)])
        got = batch.run(jobs, stream)                             # warm-up + frame count
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        batch.run(jobs, stream)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        one = digest(d_out[:out_bytes], torch)

        d_out.zero_()
        blocks = shard.split_payload(cfg, in_bytes, args.blocks, align_frames=16)
        items, at = [], 0
        for blk in blocks:
            nb = blk.frame_count * C * bps
            # every block job points at ITS slice of the payload only (a view of the resident buffer)
            items.append(dict(d_in=d_in.data_ptr() + blk.in_offset, in_offset=blk.in_offset, in_bytes=blk.in_bytes,
                              payload_bytes=in_bytes, first_frame=blk.first_frame, frame_count=blk.frame_count,
                              d_out=d_out.data_ptr() + at, out_cap_bytes=nb + 16, dither_seed=7))
            at += nb
        a.record()
        batch.run_blocks(items, stream)
        b.record()
        torch.cuda.synchronize()
        ms_blocks = a.elapsed_time(b)
        split = digest(d_out[:out_bytes], torch)
        st = batch.stats(stream)
        batch.close()
        print(json.dumps({"config": name, "what": what, "minutes": args.minutes, "in_gb": in_bytes / 1e9,
                          "out_gb": out_bytes / 1e9, "frames_per_channel": frames, "frames_ok": got[0] == frames,
                          "kernel_ms": ms, "realtime_factor": args.minutes * 60 / (ms / 1e3),
                          "dsd_samples_per_s": in_bytes * 8 / (ms / 1e3), "blocks": args.blocks,
                          "blocks_ms": ms_blocks, "digest_one_pass": one, "digest_blocks": split,
                          "blocks_equal_one_pass": one == split, "peak_abs": st["peak_abs"],
                          "plan": {k: info[k] for k in ("decim_ratio", "upsample_ratio", "num_stages", "tables")}}), flush=True)
        del d_in, d_out
        torch.cuda.empty_cache()


def shard_main(args):
    """SURVEY 8e on real GPUs: time blocks of ONE long payload spread over the ranks, no collective on the data path
    (NCCL only gathers the 64-bit digests)."""
    import torch
    import torch.distributed as dist
    import dsd2dxd_b200 as d2p
    from dsd2dxd_b200 import shard

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for name in args.only.split(","):
        kw = dict(CONFIGS[name])
        what = kw.pop("what")
        C = kw["channels"]
        per_ch = int(args.minutes * 60 * 2822400 * kw["dsd_rate"] / 8) // 4096 * 4096
        in_bytes = per_ch * C
        cfg = d2p.make_config(block_size=4096, dither_seed=7, filt="E", device=local, **kw)
        bps = d2p.plan(cfg)["bytes_per_sample"]
        blocks = shard.split_payload(cfg, in_bytes, args.blocks, align_frames=16)
        mine = shard.blocks_of_rank(blocks, rank, world)
        gen = torch.Generator(device=dev)
        gen.manual_seed(0x5D2DDD)
        # the "file": every rank can read any part of it; a rank only ever touches the slices of its own blocks
        d_in = torch.empty(in_bytes, dtype=torch.uint8, device=dev)
        for off in range(0, in_bytes, 1 << 30):
            n = min(1 << 30, in_bytes - off)
            d_in[off:off + n] = torch.randint(0, 256, (n,), dtype=torch.uint8, device=dev, generator=gen)
        batch = d2p.Batch(cfg)
        outs = [torch.zeros(b.frame_count * C * bps + 16, dtype=torch.uint8, device=dev) for b in mine]
        slices = [d_in[b.in_offset:b.in_offset + b.in_bytes].clone() for b in mine]      # the rank's private copies
        dist.barrier()
        torch.cuda.synchronize()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        batch.run_blocks([dict(d_in=s_.data_ptr(), in_offset=b.in_offset, in_bytes=b.in_bytes, payload_bytes=in_bytes,
                               first_frame=b.first_frame, frame_count=b.frame_count, d_out=o.data_ptr(),
                               out_cap_bytes=o.numel(), dither_seed=7) for b, s_, o in zip(mine, slices, outs)], stream)
        e.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(e)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        digs = {b.index: digest(o[:b.frame_count * C * bps], torch) for b, o in zip(mine, outs)}
        gathered = [None] * world
        dist.all_gather_object(gathered, digs)
        if rank == 0:
            frames = d2p.frames_for_bytes(cfg, per_ch)
            whole = torch.zeros(frames * C * bps + 16, dtype=torch.uint8, device=dev)
            jobs = batch.make_jobs([dict(d_in=d_in.data_ptr(), in_bytes=in_bytes, d_out=whole.data_ptr(),
                                         out_cap_bytes=whole.numel(), dither_seed=7)])
            batch.run(jobs, stream)
            torch.cuda.synchronize()
            want = {b.index: digest(whole[b.first_frame * C * bps:(b.first_frame + b.frame_count) * C * bps], torch)
                    for b in blocks}
            got = {k: v for g in gathered for k, v in g.items()}
            print(json.dumps({"config": name, "what": what, "minutes": args.minutes, "ranks": world, "blocks": len(blocks),
                              "sharded_ms_max_over_ranks": float(t[0]), "dsd_samples_per_s": in_bytes * 8 / (float(t[0]) / 1e3),
                              "blocks_equal_one_pass": got == want, "mismatching_blocks": [k for k in want if got.get(k) != want[k]]}),
                  flush=True)
            del whole
        batch.close()
        del d_in, outs, slices
        torch.cuda.empty_cache()
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
