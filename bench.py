#!/usr/bin/env python3
"""bench.py -- throughput of the DSD->PCM hot path on B200 (one process per GPU).

Workload = BASELINE.json configs[4] as written: a FIXED batch of 512 synthetic 10-minute DSD256
stereo files (DSF layout: planar 4096-byte blocks, LSB-first, uniform random bytes = worst case for LUT bank conflicts)
-> 24-bit 352.8 kHz PCM with the reference CLI's defaults (-r 352800 -b 24 -t E, dither T), sharded over the ranks with
shard.assign_files (strong scaling: the batch does not grow with the GPU count; no collective on the data path).
One step = one pass over the whole batch; a rank converts its files in waves of --wave files resident in HBM (the batch
is 867 GB), one launch per wave.

  value        : DSD samples/s (bits per channel x channels), whole job, inputs resident in HBM, CUDA-event timed
  e2e          : the same metric through the C-ABI host call (d2p_batch_convert_host) with pinned HOST buffers, H2D +
                 kernels + D2H inside the timed region, on --e2e-total-files files per step shared by the ranks;
                 e2e.platform_ceiling = the same bytes as bare copies issued by ALL ranks at once
  roofline     : the unit that binds the kernel -- shared-memory LUT look-ups -- against the look-up rate measured on
                 this GPU in this run (tools/microbench); roofline_hbm: algorithmic HBM bytes against MEASURED_PEAKS.json
  paths        : the other BASELINE configs' kernels (3a, 3b, 4, 4q, 1) timed in the same run, same clocks
  parity       : file 0 of both timed paths against the oracle on a 30 s prefix, with the configured dither
  cpu_baseline : the reference algorithm restated in C++ (oracle/), one file per thread, bounded sample, at nproc and at
                 the reference's own default of nproc/2 threads (N=1 only)

`--impl reference` times only that CPU arm (rank 0), as the reference Rust binary cannot be built here.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DSD256_BYTES_PER_SEC = 11289600 // 8          # per channel
SM_COUNT, LDS_LOOKUPS_PER_CLK_SM = 148, 16    # 128 B/clk/SM shared-memory crossbar / 8-byte entries


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--total-files", type=int, default=512, help="files in the batch (BASELINE config 5), sharded over the ranks")
    ap.add_argument("--wave", type=int, default=8, help="files resident in HBM per launch on one GPU (867 GB do not fit at once)")
    ap.add_argument("--minutes", type=float, default=10.0, help="duration of each synthetic file")
    ap.add_argument("--dither", default="T", choices=["T", "R", "X"], help="CLI default for 24-bit is T")
    ap.add_argument("--e2e-total-files", type=int, default=16, help="files per step on the host-buffer (e2e) path, all ranks together")
    ap.add_argument("--cpu-seconds", type=float, default=30.0, help="audio seconds per file in the CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-paths", action="store_true", help="skip the per-path timings (configs 3a, 3b, 4, 4q)")
    ap.add_argument("--path-seconds", type=float, default=120.0, help="audio seconds of each per-path workload")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def measure_lds_peak(local):
    """Conflict-free LDS.64 look-up rate of THIS GPU under THIS run's clocks (tools/microbench.cu, < 1 s): the
    denominator of the look-up roofline.  Falls back to the committed round-1 measurement, then to the nominal figure."""
    exe = os.path.join(ROOT, "tools", "microbench")
    try:
        env = dict(os.environ)
        vis = env.get("CUDA_VISIBLE_DEVICES")
        ids = [x.strip() for x in vis.split(",")] if vis else None
        env["CUDA_VISIBLE_DEVICES"] = ids[local] if ids and local < len(ids) else str(local)
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120, env=env).stdout
        mb = json.loads(out.strip().splitlines()[-1])
        return float(mb["lds64_lookups_per_s"]), "measured in this run: tools/microbench.cu (conflict-free LDS.64 look-ups) on this GPU", mb
    except Exception:
        pass
    try:
        mb = json.load(open(os.path.join(ROOT, "profiles", "r01_microbench.json")))
        return float(mb["lds64_lookups_per_s"]), "measured earlier on this pool: profiles/r01_microbench.json", mb
    except Exception:
        pk, _ = peaks()
        return SM_COUNT * LDS_LOOKUPS_PER_CLK_SM * pk.get("sm_max_mhz", 1965.0) * 1e6, "nominal 16 lookups/clk/SM x 148 SM x max SM clock", None


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(r[1]) for r in self.rows if len(r) >= 8 and r[1].isdigit())
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        mx = [int(r[2]) for r in self.rows if len(r) >= 8 and r[2].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_run(args, threads):
    """Reference algorithm restated in C++ (oracle), one file per thread, bounded sample."""
    import numpy as np
    from oracle import orclib
    per_ch = int(args.cpu_seconds * DSD256_BYTES_PER_SEC) // 4096 * 4096
    raw = np.random.default_rng(0x5D2DDD).integers(0, 256, size=2 * per_ch, dtype=np.uint8)
    nfiles = threads
    secs = orclib.bench(raw, nfiles, threads, dsd_rate=4, out_rate=352800, filt="E", bits=24, dither=args.dither,
                        channels=2, layout="P", lsb_first=True, block_size=4096)
    samples = nfiles * 2 * per_ch * 8
    return {"value": samples / secs, "unit": "DSD samples/s", "cores": threads, "seconds": secs,
            "realtime_factor": nfiles * args.cpu_seconds / secs}


def cpu_baseline(args):
    """Both thread policies: every host core, and the reference's own default of available_parallelism() / 2
    (src/main.rs:148-163).  The headline `value` is the faster one of the two."""
    ncpu = os.cpu_count() or 1
    full, half = cpu_run(args, ncpu), cpu_run(args, max(1, ncpu // 2))
    best = full if full["value"] >= half["value"] else half
    out = dict(best)
    out.update({"kind": "port",
                "sample": f"one file per thread x {args.cpu_seconds:g} s of DSD256 stereo each, DSP only (reference algorithm "
                          f"restated in C++, oracle/; the Rust binary cannot be built on this box: no cargo, dsd_reader missing)",
                "threads_all_cores": full, "threads_reference_default": half})
    return out


def run_reference(args, rank, world):
    if rank != 0:
        return
    vals = []
    t0 = time.perf_counter()
    for _ in range(max(1, args.steps)):
        vals.append(cpu_baseline(args))
    wall = time.perf_counter() - t0
    v = sum(x["value"] for x in vals) / len(vals)
    cb = dict(vals[-1]); cb["value"] = v
    line = {"impl": "reference", "metric": "dsd_samples_per_sec", "value": v, "unit": "DSD samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000 * wall / max(1, args.steps),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world), "cpu_baseline": cb,
            "realtime_factor": cb["realtime_factor"],
            "e2e": {"value": v, "unit": "DSD samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(args, world):
    return {"workload": f"BASELINE config 5: batch of {args.total_files} synthetic {args.minutes:g}-minute DSD256 stereo files "
                        f"(DSF layout: planar 4096-byte blocks, LSB-first, random bytes) -> 24-bit 352.8 kHz, filter E "
                        f"(HTAPS_DSD256_32TO1_EQ, 151 half taps, decim 32), dither {args.dither}; files assigned to the ranks "
                        f"with shard.assign_files and converted in waves of {args.wave} resident files per launch",
            "total_files": args.total_files, "wave_files": args.wave, "minutes_per_file": args.minutes, "dsd_rate": "DSD256",
            "channels": 2, "out_rate": 352800, "bits": 24, "dither": args.dither,
            "l2": "every launch reads and writes tens of GB: far larger than L2 (no flush needed)"}


# ---- per-path timings: every kernel gets a clock-stamped number from the same run ----------------
PATHS = [
    ("3a", "DSD256 stereo planar -> 20-bit 176.4 kHz (decim 64, HTAPS_DSD256_64TO1_EQ)", dict(dsd_rate=4, out_rate=176400, bits=20, channels=2)),
    ("3b", "DSD256 stereo planar -> f32 192 kHz (L/M 5/294, two stages)", dict(dsd_rate=4, out_rate=192000, bits=32, channels=2)),
    ("3b-96k", "DSD256 stereo planar -> f32 96 kHz (L/M 5/588, h2 = 173)", dict(dsd_rate=4, out_rate=96000, bits=32, channels=2)),
    ("4", "DSD512 6-ch byte-interleaved MSB-first -> 24-bit 352.8 kHz (decim 64, HTAPS_DSD512_64TO1_EQ)",
     dict(dsd_rate=8, out_rate=352800, bits=24, channels=6, layout="I", lsb_first=False)),
    ("4q", "DSD512 6-ch byte-interleaved MSB-first -> 24-bit '384 kHz' as written (L/M quirk path, SURVEY F8: 768 k frames/s)",
     dict(dsd_rate=8, out_rate=384000, bits=24, channels=6, layout="I", lsb_first=False)),
    ("1", "DSD64 stereo planar -> 24-bit 352.8 kHz (decim 8, HTAPS_DSD64_8TO1_EQ)", dict(dsd_rate=1, out_rate=352800, bits=24, channels=2)),
]


def smem_reads(info, in_bytes_per_ch, channels, frames):
    """Algorithmic 8-byte shared-memory reads of one conversion: 2*nt per frame on the integer path;
    groups per stage-1 output + 2*h per stage-2/3 output on the L/M path."""
    if info["num_stages"] == 1:
        nt = (info["half_taps"][0] + 7) // 8
        return channels * in_bytes_per_ch * 2 * nt / (info["decims"][0] // 8)
    L, h1 = info["upsample_ratio"], info["half_taps"][0]
    cdiv = lambda a, b: -(-a // b)
    groups = [cdiv(cdiv(2 * h1 - p, L), 8) for p in range(L)]               # 8-tap look-up groups of phase p
    n_last = frames
    reads = 0.0
    for st in range(info["num_stages"] - 1, 0, -1):                        # stage 3 / 2: 2h reads per output
        reads += n_last * 2 * info["half_taps"][st]
        n_last = n_last * info["decims"][st]
    reads += n_last * sum(groups) / L                                       # stage 1: one look-up per group
    return channels * reads


def time_paths(args, d2p, torch, dev, lds_peak, check=True):
    import numpy as np
    out = []
    stream = torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device=dev)
    gen.manual_seed(0x5D2DDD)
    for name, what, kw in PATHS:
        rate, ch = kw["dsd_rate"], kw["channels"]
        per_ch = int(args.path_seconds * 2822400 * rate / 8) // 4096 * 4096
        cfg = d2p.make_config(dither="X", block_size=4096, dither_seed=1, device=dev.index, **kw)
        info = d2p.plan(cfg)
        frames = d2p.frames_for_bytes(cfg, per_ch)
        d_in = torch.randint(0, 256, (per_ch * ch,), dtype=torch.uint8, device=dev, generator=gen)
        d_out = torch.empty(frames * ch * info["bytes_per_sample"] + 16, dtype=torch.uint8, device=dev)
        batch = d2p.Batch(cfg)
        jobs = batch.make_jobs([dict(d_in=d_in.data_ptr(), in_bytes=d_in.numel(), d_out=d_out.data_ptr(), out_cap_bytes=d_out.numel())])
        for _ in range(2):
            batch.run(jobs, stream)
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); batch.run(jobs, stream); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        batch.close()
        ms = sorted(ts)[len(ts) // 2]
        # the timed output against the oracle on the first seconds (causal filters: a prefix of the input gives a prefix
        # of the output; block groups / interleaved frames both put the first pre * ch bytes of the payload first)
        prefix_equal, pre_s = None, min(2.0, args.path_seconds)
        if check:
            from oracle import orclib
            pre = int(pre_s * 2822400 * rate / 8) // 4096 * 4096
            ref = orclib.convert(d_in[:pre * ch].cpu().numpy(), want_f64=False, dither="X", **kw)
            prefix_equal = bool(ref.pcm.size > 0 and np.array_equal(d_out[:ref.pcm.size].cpu().numpy(), ref.pcm))
        reads = smem_reads(info, per_ch, ch, frames)
        out.append({"config": name, "workload": f"{args.path_seconds:g} s of {what}, device-resident, dither X",
                    "kernel": "d2p::lutfir_kernel" if info["num_stages"] == 1 else "d2p::lm_kernel", "kernel_ms": ms,
                    "dsd_samples_per_s": per_ch * 8 * ch / (ms / 1e3), "realtime_factor": args.path_seconds / (ms / 1e3),
                    "smem_reads_8B": reads, "lds_frac": reads / (ms / 1e3) / lds_peak, "seconds": args.path_seconds, "dither": "X",
                    "oracle_prefix_equal": prefix_equal, "oracle_prefix_seconds": pre_s if check else None})
        del d_in, d_out
    return out


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    import dsd2dxd_b200 as d2p
    from dsd2dxd_b200 import shard

    if d2p.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the library has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly one JSON line: NCCL's own banner ("NCCL version ...") is sent to stderr instead
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    # look-up roofline denominator, measured on rank 0's GPU while the others wait
    lds_peak, lds_kind, mb = (measure_lds_peak(local) if rank == 0 else (0.0, "", None))
    barrier()

    per_ch = int(args.minutes * 60 * DSD256_BYTES_PER_SEC) // 4096 * 4096
    in_bytes = 2 * per_ch
    cfg = d2p.make_config(dsd_rate=4, out_rate=352800, filt="E", bits=24, dither=args.dither, channels=2,
                          layout="P", lsb_first=True, block_size=4096, device=local, dither_seed=0x5D2DDD)
    info = d2p.plan(cfg)
    frames = d2p.frames_for_bytes(cfg, per_ch)
    out_bytes = frames * 2 * 3
    batch = d2p.Batch(cfg)

    # BASELINE config 5: a FIXED list of files, balanced over the ranks (strong scaling); every rank converts its share
    # in waves of `wave` device-resident files.  The synthetic payload of file i is buffer i % wave (generated on the
    # device from seed 0x5D2DDD + buffer index); its dither seed is 0x5D2DDD + i.
    mine = shard.assign_files([in_bytes] * args.total_files, world)[rank]
    wave = max(1, min(args.wave, max(1, len(mine))))
    gen = torch.Generator(device=dev)
    d_in, d_out = [], []
    for f in range(wave):
        gen.manual_seed(0x5D2DDD + f)
        d_in.append(torch.randint(0, 256, (in_bytes,), dtype=torch.uint8, device=dev, generator=gen))
        d_out.append(torch.empty(out_bytes + 16, dtype=torch.uint8, device=dev))
    waves = []
    for w0 in range(0, len(mine), wave):
        part = mine[w0:w0 + wave]
        waves.append(batch.make_jobs(dict(d_in=d_in[k].data_ptr(), in_bytes=in_bytes, d_out=d_out[k].data_ptr(),
                                          out_cap_bytes=d_out[k].numel(), dither_seed=0x5D2DDD + g) for k, g in enumerate(part)))
    stream = torch.cuda.current_stream().cuda_stream

    def one_step():
        for jobs in waves:
            batch.run(jobs, stream)

    for _ in range(args.warmup):
        one_step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = batch.launches()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for s in range(args.steps):
        one_step()
        ev[s + 1].record()
    barrier()
    total_ms = ev[0].elapsed_time(ev[args.steps])
    launches = batch.launches() - launches0
    clocks = sampler.stop()
    # the dominant kernel, timed alone on its stream: one launch of a full wave
    ka, kb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kts = []
    for _ in range(3):
        ka.record(); batch.run(waves[0], stream); kb.record()
        torch.cuda.synchronize()
        kts.append(ka.elapsed_time(kb))
    kernel_ms = sorted(kts)[1]
    kernel_files = len(waves[0])

    # ---- e2e: host buffers through the C ABI, copies inside the timed region -----------------------------------------
    e2e_mine = shard.assign_files([in_bytes] * args.e2e_total_files, world)[rank]
    h_in = d2p.HostBuffer(in_bytes)
    h_out = d2p.HostBuffer(out_bytes)
    h_in.array[:] = d_in[0].cpu().numpy()
    got = batch.convert_host(h_in.array, h_out.array)                # warm: staging buffers, streams
    barrier()
    t0 = time.perf_counter()
    for s in range(args.steps):
        for _ in e2e_mine:
            got = batch.convert_host(h_in.array, h_out.array)
    torch.cuda.synchronize()
    e2e_s = allmax(time.perf_counter() - t0)
    assert got == frames
    e2e_pcm = h_out.array[:out_bytes].copy()

    # ---- what the box gives bare pinned copies when EVERY rank copies at once (the ceiling of any host-buffer path) --
    th_in, th_out = torch.from_numpy(h_in.array), torch.from_numpy(h_out.array)
    cs = [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)]

    e2e_files_rank = max(1, max(len(x) for x in shard.assign_files([1] * args.e2e_total_files, world)))

    def timed_copies(do_h2d, do_d2h, reps=3):
        """Seconds per file of `e2e_files_rank` files' copies issued back to back (H2D on one stream, D2H on another, as
        the e2e loop keeps them in flight across files): a pipelined conversion cannot be faster than this."""
        best = None
        for _ in range(reps):
            barrier()
            t1 = time.perf_counter()
            for _f in range(e2e_files_rank):
                if do_h2d:
                    with torch.cuda.stream(cs[0]):
                        d_in[0].copy_(th_in, non_blocking=True)
                if do_d2h:
                    with torch.cuda.stream(cs[1]):
                        th_out.copy_(d_out[0][:out_bytes], non_blocking=True)
            torch.cuda.synchronize()
            dt = allmax(time.perf_counter() - t1) / e2e_files_rank    # slowest rank
            best = dt if best is None else min(best, dt)
        return best

    t_h2d, t_d2h, t_both = timed_copies(True, False), timed_copies(False, True), timed_copies(True, True)
    e2e_file_s = e2e_s / (args.steps * e2e_files_rank)                # per file on the busiest rank
    ceiling = {"what": "bare pinned cudaMemcpyAsync of the payloads (H2D) and PCM (D2H) of as many files per rank as one e2e step "
                       "converts, issued back to back on two streams, ALL ranks at once behind a barrier, slowest rank, best "
                       "of 3, per file; GB/s are aggregates over the ranks",
               "files_per_rank": e2e_files_rank,
               "h2d_only_gbs": world * in_bytes / t_h2d / 1e9, "d2h_only_gbs": world * out_bytes / t_d2h / 1e9,
               "both_h2d_gbs": world * in_bytes / t_both / 1e9, "both_d2h_gbs": world * out_bytes / t_both / 1e9,
               "both_seconds_per_file": t_both, "e2e_seconds_per_file": e2e_file_s, "e2e_over_ceiling": e2e_file_s / t_both}

    # ---- parity of the timed paths against the oracle (rank 0, a prefix the CPU finishes in about a second) ----------
    parity = None
    if rank == 0:
        from oracle import orclib
        jobs0 = batch.make_jobs([dict(d_in=d_in[0].data_ptr(), in_bytes=in_bytes, d_out=d_out[0].data_ptr(),
                                      out_cap_bytes=d_out[0].numel(), dither_seed=0x5D2DDD)])
        batch.run(jobs0, stream)                                      # file 0 again with the seed the host path uses
        torch.cuda.synchronize()
        dev_pcm = d_out[0][:out_bytes].cpu().numpy()
        pre_s = min(30.0, args.minutes * 60)
        pre_groups = int(pre_s * DSD256_BYTES_PER_SEC) // 4096
        ref = orclib.convert(h_in.array[:pre_groups * 8192], want_f64=False, dsd_rate=4, out_rate=352800, filt="E", bits=24,
                             dither=args.dither, channels=2, dither_seed=0x5D2DDD)
        n = ref.pcm.size
        parity = {"oracle_prefix_equal": bool(np.array_equal(dev_pcm[:n], ref.pcm) and np.array_equal(e2e_pcm[:n], ref.pcm)),
                  "device_path_equal": bool(np.array_equal(dev_pcm[:n], ref.pcm)), "e2e_path_equal": bool(np.array_equal(e2e_pcm[:n], ref.pcm)),
                  "prefix_seconds": pre_s, "prefix_frames": int(ref.frames[0]), "dither": args.dither,
                  "what": "file 0 of the device-resident path and of the host-buffer path vs oracle/ on the first seconds (causal filter: a prefix of the input gives a prefix of the output)",
                  "e2e_matches_device_path": bool(np.array_equal(dev_pcm, e2e_pcm))}

    total_ms = allmax(total_ms)
    file_samples = 2 * per_ch * 8                                    # DSD samples = bits per channel x channels
    value = args.total_files * file_samples * args.steps / (total_ms / 1e3)
    e2e_value = args.e2e_total_files * file_samples * args.steps / e2e_s
    pk, pk_kind = peaks()
    nt = (info["half_taps"][0] + 7) // 8
    lookups = kernel_files * in_bytes * (2 * nt) / (info["decims"][0] // 8)
    alg_bytes = kernel_files * (in_bytes + out_bytes)                # (1/8 + 3/32) B per DSD sample
    traffic = None
    try:        # dram__bytes_read.sum + dram__bytes_write.sum of one such launch (ncu --set full of this command)
        tr = json.load(open(os.path.join(ROOT, "profiles", "r02_bench_traffic.json")))
        if tr.get("wave_files") == kernel_files and tr.get("minutes_per_file") == args.minutes and tr.get("dither") == args.dither:
            traffic = tr["dram_bytes_per_launch"]
    except Exception:
        pass

    if rank == 0:
        sm_mhz = clocks["sm_mhz"] or pk.get("sm_max_mhz", 1965.0)
        lds = {"bound": "smem_lut", "what": "shared-memory LUT look-ups (8 B each): the unit that binds this path (SURVEY 8d)",
               "achieved": lookups * 8 / (kernel_ms / 1e3) / 1e9, "peak": lds_peak * 8 / 1e9, "unit": "GB/s",
               "frac": lookups / (kernel_ms / 1e3) / lds_peak, "traffic": traffic,
               "traffic_what": "DRAM bytes of one such launch (dram__bytes_read.sum + dram__bytes_write.sum, committed ncu capture of "
                               "this command: profiles/r02_bench_traffic.json; not re-measured in this run); algorithmic HBM bytes "
                               "are in roofline_hbm",
               "peak_kind": lds_kind,
               "lookups_per_launch": lookups, "lookups_per_s": lookups / (kernel_ms / 1e3), "peak_lookups_per_s": lds_peak,
               "kernel_ms": kernel_ms, "files_per_launch": kernel_files,
               "frac_of_nominal_at_observed_clock": lookups / (kernel_ms / 1e3) / (SM_COUNT * LDS_LOOKUPS_PER_CLK_SM * sm_mhz * 1e6)}
        line = {
            "metric": "dsd_samples_per_sec", "value": value, "unit": "DSD samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "realtime_factor": value / (2 * 11289600.0),
            "e2e": {"value": e2e_value, "unit": "DSD samples/s", "h2d_bytes_per_step": args.e2e_total_files * in_bytes,
                    "d2h_bytes_per_step": args.e2e_total_files * out_bytes, "files_per_step": args.e2e_total_files,
                    "timer": "wall clock around the synchronous C-ABI calls (d2p_batch_convert_host, pinned host buffers), max over ranks",
                    "realtime_factor": e2e_value / (2 * 11289600.0), "platform_ceiling": ceiling,
                    # the same as scalars (a parser that drops nested objects keeps these)
                    "ceiling_both_s_per_file": ceiling["both_seconds_per_file"], "over_ceiling": ceiling["e2e_over_ceiling"],
                    "ceiling_h2d_gbs": ceiling["both_h2d_gbs"], "ceiling_d2h_gbs": ceiling["both_d2h_gbs"]},
            "gpu_launches": int(launches),
            "kernel": "d2p::lutfir_kernel<1,'T',3,1>" if args.dither == "T" else "d2p::lutfir_kernel", "kernel_ms": kernel_ms,
            "roofline": lds, "lds": {"frac": lds["frac"], "achieved": lds["lookups_per_s"], "peak": lds_peak, "unit": "lookups/s",
                                       "note": "same object as roofline, in look-ups per second (the round-1 key)"},
            "roofline_hbm": {"bound": "hbm", "achieved": alg_bytes / (kernel_ms / 1e3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": alg_bytes / (kernel_ms / 1e3) / 1e9 / pk["hbm_gbs"], "traffic": traffic,
                             "algorithmic_bytes": alg_bytes, "peak_kind": pk_kind,
                             "note": "HBM is not the binding unit of this path: see roofline"},
            "microbench": mb, "parity": parity, "clocks": clocks,
        }
        if not args.no_paths and world == 1:
            line["paths"] = time_paths(args, d2p, torch, dev, lds_peak)
        if not args.no_cpu and world == 1:           # the CPU leg is reported at N=1 only
            line["cpu_baseline"] = cpu_baseline(args)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
