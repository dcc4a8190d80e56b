#!/usr/bin/env python3
"""bench.py -- throughput of the DSD->PCM hot path on B200 (one process per GPU).

Workload (BASELINE.json configs[4]): a batch of synthetic 10-minute DSD256 stereo
files (DSF layout: planar 4096-byte blocks, LSB-first, uniform random bytes = worst case for LUT bank
conflicts) -> 24-bit 352.8 kHz PCM with the reference CLI's defaults (-r 352800 -b 24 -t E, dither T).
One step = one pass of the hot path over `--files` such files per GPU (weak scaling: every rank
converts its own shard, no collective on the data path).

  value : DSD samples/s (bits per channel x channels), whole job, inputs resident in HBM, CUDA-event timed
  e2e   : the same metric through the C-ABI host call (d2p_batch_convert_host) with pinned HOST buffers,
          H2D + kernels + D2H inside the timed region (wall clock around the synchronous call)
  roofline / lds : algorithmic HBM bytes and shared-memory LUT look-ups per launch over the measured
          kernel time (DESIGN.md "Measurement")
  cpu_baseline : the reference algorithm restated in C++ (oracle/), one file per thread, on a bounded sample

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
    ap.add_argument("--files", type=int, default=8, help="10-minute DSD256 stereo files per GPU per step")
    ap.add_argument("--minutes", type=float, default=10.0, help="duration of each synthetic file")
    ap.add_argument("--dither", default="T", choices=["T", "R", "X"], help="CLI default for 24-bit is T")
    ap.add_argument("--e2e-files", type=int, default=2, help="files per step on the host-buffer (e2e) path")
    ap.add_argument("--cpu-seconds", type=float, default=30.0, help="audio seconds per file in the CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


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


def cpu_baseline(args, cores):
    """Reference algorithm restated in C++ (oracle), one file per thread, bounded sample."""
    import numpy as np
    from oracle import orclib
    per_ch = int(args.cpu_seconds * DSD256_BYTES_PER_SEC) // 4096 * 4096
    raw = np.random.default_rng(0x5D2DDD).integers(0, 256, size=2 * per_ch, dtype=np.uint8)
    nfiles = cores
    secs = orclib.bench(raw, nfiles, cores, dsd_rate=4, out_rate=352800, filt="E", bits=24, dither=args.dither,
                        channels=2, layout="P", lsb_first=True, block_size=4096)
    samples = nfiles * 2 * per_ch * 8
    return {"value": samples / secs, "unit": "DSD samples/s", "cores": cores, "kind": "port",
            "sample": f"{nfiles} files x {args.cpu_seconds:g} s of DSD256 stereo, one file per thread, DSP only "
                      f"(reference algorithm restated in C++; the Rust binary cannot be built on this box)",
            "seconds": secs, "realtime_factor": nfiles * args.cpu_seconds / secs}


def run_reference(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    vals = []
    for _ in range(args.warmup):
        pass                                  # nothing to warm on the CPU arm beyond the first timed step
    t0 = time.perf_counter()
    for _ in range(max(1, args.steps)):
        vals.append(cpu_baseline(args, cores))
    wall = time.perf_counter() - t0
    v = sum(x["value"] for x in vals) / len(vals)
    cb = dict(vals[-1]); cb["value"] = v
    line = {"impl": "reference", "metric": "dsd_samples_per_sec", "value": v, "unit": "DSD samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000 * wall / max(1, args.steps),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args), "cpu_baseline": cb,
            "realtime_factor": cb["realtime_factor"],
            "e2e": {"value": v, "unit": "DSD samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(args):
    return {"workload": f"batch of {args.files} synthetic {args.minutes:g}-minute DSD256 stereo files per GPU "
                        f"(DSF layout: planar 4096-byte blocks, LSB-first, random bytes) -> 24-bit 352.8 kHz, "
                        f"filter E (HTAPS_DSD256_32TO1_EQ, 151 half taps, decim 32), dither {args.dither}",
            "files_per_gpu": args.files, "minutes_per_file": args.minutes, "dsd_rate": "DSD256", "channels": 2,
            "out_rate": 352800, "bits": 24, "dither": args.dither,
            "l2": "inputs+outputs per step are far larger than L2 (no flush needed)"}


_FULL_AFFINITY = None


def bind_to_gpu_numa(local):
    """One process per GPU: run this rank's host thread (and, by first touch, its pinned staging memory) on the CPUs
    NVML reports as local to the GPU.  Returns a short description for the JSON line; never fatal."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else local
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            global _FULL_AFFINITY
            _FULL_AFFINITY = os.sched_getaffinity(0)
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} cpus local to gpu {idx}"
    except Exception as e:      # noqa: BLE001
        return f"unbound ({type(e).__name__})"
    return "unbound"


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

    if d2p.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the library has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa(local)          # pinned host buffers of the e2e path then live next to this rank's GPU
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

    per_ch = int(args.minutes * 60 * DSD256_BYTES_PER_SEC) // 4096 * 4096
    in_bytes = 2 * per_ch
    cfg = d2p.make_config(dsd_rate=4, out_rate=352800, filt="E", bits=24, dither=args.dither, channels=2,
                          layout="P", lsb_first=True, block_size=4096, device=local, dither_seed=0x5D2DDD)
    info = d2p.plan(cfg)
    frames = d2p.frames_for_bytes(cfg, per_ch)
    out_bytes = frames * 2 * 3
    batch = d2p.Batch(cfg)

    # synthetic inputs generated on the device (seed 0x5D2DDD + global file index), outputs preallocated
    gen = torch.Generator(device=dev)
    d_in, d_out = [], []
    for f in range(args.files):
        gen.manual_seed(0x5D2DDD + rank * args.files + f)
        d_in.append(torch.randint(0, 256, (in_bytes,), dtype=torch.uint8, device=dev, generator=gen))
        d_out.append(torch.empty(out_bytes + 16, dtype=torch.uint8, device=dev))
    jobs = batch.make_jobs(dict(d_in=i.data_ptr(), in_bytes=in_bytes, d_out=o.data_ptr(), out_cap_bytes=o.numel(),
                                dither_seed=0x5D2DDD + rank * args.files + k)
                           for k, (i, o) in enumerate(zip(d_in, d_out)))
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        batch.run(jobs, stream)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = batch.launches()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for s in range(args.steps):
        batch.run(jobs, stream)
        ev[s + 1].record()
    barrier()
    step_ms = [ev[s].elapsed_time(ev[s + 1]) for s in range(args.steps)]
    total_ms = ev[0].elapsed_time(ev[args.steps])
    launches = batch.launches() - launches0

    # ---- e2e: host buffers through the C ABI, copies inside the timed region ----
    nf = max(1, min(args.e2e_files, args.files))
    h_in = d2p.HostBuffer(in_bytes)
    h_out = d2p.HostBuffer(out_bytes)
    h_in.array[:] = d_in[0].cpu().numpy()
    batch.convert_host(h_in.array, h_out.array)                     # warm: staging buffers, streams
    barrier()
    t0 = time.perf_counter()
    for s in range(args.steps):
        for _ in range(nf):
            got = batch.convert_host(h_in.array, h_out.array)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert got == frames
    # what the same bytes cost as bare copies (H2D of one file's payload and D2H of its PCM at once, pinned memory, two
    # streams): the ceiling of any host-buffer path on this box, reported beside e2e
    th_in, th_out = torch.from_numpy(h_in.array), torch.from_numpy(h_out.array)
    cs = [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)]
    copy_s = []
    for _ in range(3):
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        with torch.cuda.stream(cs[0]):
            d_in[0].copy_(th_in, non_blocking=True)
        with torch.cuda.stream(cs[1]):
            th_out.copy_(d_out[0][:out_bytes], non_blocking=True)
        torch.cuda.synchronize()
        copy_s.append(time.perf_counter() - t1)
    copy_file_s = min(copy_s)
    # the host path must give the same bytes as the device-resident path
    same = bool(np.array_equal(h_out.array[:out_bytes], d_out[0][:out_bytes].cpu().numpy())) if args.dither == "X" else None
    clocks = sampler.stop()

    t = torch.tensor([total_ms, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_s = float(t[0]), float(t[1])

    samples_per_step = args.files * 2 * per_ch * 8 * world          # DSD samples = bits per channel x channels
    value = samples_per_step * args.steps / (total_ms / 1e3)
    e2e_value = nf * 2 * per_ch * 8 * world * args.steps / e2e_s
    kernel_ms = sum(step_ms) / len(step_ms)                          # one kernel launch per step on this stream
    pk, pk_kind = peaks()
    alg_bytes = args.files * (in_bytes + out_bytes)                  # (1/8 + 3/32) B per DSD sample
    lookups = args.files * in_bytes * (2 * ((info["half_taps"][0] + 7) // 8)) / (info["decims"][0] // 8)
    sm_mhz = clocks["sm_mhz"] or pk.get("sm_max_mhz", 1965.0)
    lds_peak = SM_COUNT * LDS_LOOKUPS_PER_CLK_SM * pk.get("sm_max_mhz", 1965.0) * 1e6
    lds_kind = "nominal 16 lookups/clk/SM x 148 SM x max SM clock"
    try:        # measured on this pool's B200 by tools/microbench.cu (conflict-free LDS.64 look-ups)
        mb = json.load(open(os.path.join(ROOT, "profiles", "r01_microbench.json")))
        lds_peak = float(mb["lds64_lookups_per_s"])
        lds_kind = "measured: tools/microbench.cu on B200 (profiles/r01_microbench.json)"
    except Exception:
        pass
    traffic = None
    try:        # dram__bytes_read.sum + dram__bytes_write.sum of one launch of this very command (ncu --set full)
        tr = json.load(open(os.path.join(ROOT, "profiles", "r01_bench_traffic.json")))
        if tr.get("files_per_gpu") == args.files and tr.get("minutes_per_file") == args.minutes and tr.get("dither") == args.dither:
            traffic = tr["dram_bytes_per_launch"]
    except Exception:
        pass

    if rank == 0:
        line = {
            "metric": "dsd_samples_per_sec", "value": value, "unit": "DSD samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args),
            "realtime_factor": value / (2 * 11289600.0),
            "e2e": {"value": e2e_value, "unit": "DSD samples/s", "h2d_bytes_per_step": nf * in_bytes,
                    "d2h_bytes_per_step": nf * out_bytes, "files_per_step": nf, "timer": "wall clock around the synchronous host call",
                    "realtime_factor": e2e_value / (2 * 11289600.0), "matches_device_path": same, "host_affinity": numa,
                    "copies_alone": {"what": "bare pinned H2D of one file's payload + D2H of its PCM on two streams, rank 0",
                                     "seconds_per_file": copy_file_s, "h2d_gbs": in_bytes / copy_file_s / 1e9,
                                     "d2h_gbs": out_bytes / copy_file_s / 1e9,
                                     "e2e_over_copy_time": (e2e_s / (nf * args.steps)) / copy_file_s}},
            "gpu_launches": int(launches),
            "kernel": "d2p::lutfir_kernel", "kernel_ms": kernel_ms,
            "roofline": {"bound": "hbm", "achieved": alg_bytes / (kernel_ms / 1e3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                         "frac": alg_bytes / (kernel_ms / 1e3) / 1e9 / pk["hbm_gbs"], "traffic": traffic,
                         "algorithmic_bytes": alg_bytes, "peak_kind": pk_kind,
                         "note": "HBM is not the binding unit of this path (SURVEY.md 8d): see lds"},
            "lds": {"bound": "shared-memory LUT look-ups (8 B each)", "achieved": lookups / (kernel_ms / 1e3),
                    "peak": lds_peak, "unit": "lookups/s", "frac": lookups / (kernel_ms / 1e3) / lds_peak,
                    "peak_kind": lds_kind, "lookups_per_launch": lookups,
                    "frac_at_observed_clock": lookups / (kernel_ms / 1e3) / (SM_COUNT * LDS_LOOKUPS_PER_CLK_SM * sm_mhz * 1e6)},
            "clocks": clocks,
        }
        if not args.no_cpu and world == 1:           # the CPU leg is reported at N=1 only
            if _FULL_AFFINITY:                       # the CPU arm uses every host core, not just the GPU's neighbours
                os.sched_setaffinity(0, _FULL_AFFINITY)
            line["cpu_baseline"] = cpu_baseline(args, os.cpu_count() or 1)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
