"""The native caller of the C ABI (tools/d2p_cli.cpp) and the boundary pieces added with it:
cancellation flag, progress callback, one peak per job for dsd_levels batches, plan-time rejection of block sizes the
closed forms do not cover, overflow-proof container size checks.

CPU part: both headers compile as C99, the CLI builds against the library alone and refuses to run without a GPU,
plan / parser checks.  GPU part: CLI file -> file and stdin -> stdout (4096-byte block pushes, the reference's own
call pattern) against the oracle; cancel / progress / per-job statistics.
"""
import ctypes as C
import os
import struct
import subprocess

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import _capi
from oracle import containers as orc
from oracle import orclib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "tools", "d2p_cli")


@pytest.fixture(scope="module")
def cli():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tools"), "-s", "d2p_cli"])
    return CLI


@pytest.mark.parametrize("header", ["dsd2pcm_b200.h", "dsd2pcm_b200_files.h"])
def test_headers_are_plain_c99(header):
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-x", "c",
                           os.path.join(ROOT, "include", header)])


def test_cli_links_the_c_abi_only(cli):
    ldd = subprocess.run(["ldd", cli], capture_output=True, text=True).stdout
    assert "libdsd2pcm_b200.so" in ldd and "libpython" not in ldd
    r = subprocess.run([cli], capture_output=True, text=True)
    assert r.returncode == 2 and "usage: d2p_cli" in r.stderr
    if d2p.device_count() == 0:                      # no CPU path: the CLI says so instead of converting
        r = subprocess.run([cli, "-o", "W", "x.dsf"], capture_output=True, text=True)
        assert r.returncode == 1 and "no CUDA device" in r.stderr


def test_block_size_must_hold_whole_output_frames():
    """process_bytes restarts its frame phase per call (rd/byte_precalc_decimator.rs:121-123): a block that is not a
    multiple of bytes-per-output-frame is refused, not converted differently."""
    ok = d2p.make_config(dsd_rate=1, out_rate=88200, block_size=4096)         # bpo = 4
    assert d2p.plan(ok)["decim_ratio"] == 32
    for bad in (4098, 6, 1001):
        with pytest.raises(_capi.D2PError, match="not a multiple") as e:
            d2p.plan(d2p.make_config(dsd_rate=1, out_rate=88200, block_size=bad))
        assert e.value.status == _capi.ERR_INVALID_ARG
    d2p.plan(d2p.make_config(dsd_rate=1, out_rate=96000, block_size=1001))    # L/M path consumes bits: any block size
    with pytest.raises(_capi.D2PError, match="2\\^31"):
        d2p.plan(d2p.make_config(dsd_rate=1, out_rate=352800, block_size=1 << 31))


def test_container_sizes_that_would_wrap_are_rejected():
    pay = bytes(8192)
    dsf = bytearray(orc.make_dsf(pay, 2, 1))
    struct.pack_into("<Q", dsf, 28 + 4, (1 << 64) - 40)                      # fmt chunk size: 28 + size + 12 wraps to 0
    with pytest.raises(_capi.D2PError, match="bad fmt chunk size"):
        d2p.container_parse(bytes(dsf))
    dff = bytearray(orc.make_dff(pay, 2, 1))
    at = dff.index(b"FS  ")
    struct.pack_into(">Q", dff, at + 4, (1 << 64) - 1)                       # PROP sub-chunk size: q + 12 + size wraps
    with pytest.raises(_capi.D2PError):                                      # walk stops: no sample rate -> not a DSD file
        d2p.container_parse(bytes(dff))


# ---- GPU ---------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_cli_file_to_file_matches_oracle(cli, ref_files, tmp_path):
    src = tmp_path / "pink.dsf"
    src.write_bytes(open(ref_files["test/pinknoise_stereo_128.dsf"], "rb").read())
    r = subprocess.run([cli, "-o", "W", "-r", "176400", "-b", "24", "-d", "X", "-a", str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "DSP speed for pink.dsf:" in r.stderr and "End-to-end:" in r.stderr      # report_timing's own wording
    out = tmp_path / "pink_176_4K.wav"                                               # -a: get_out_filename_path suffix
    info = d2p.container_open(src)
    pay = np.fromfile(src, dtype=np.uint8)[info["data_offset"]: info["data_offset"] + info["data_bytes"]]
    ref = orclib.convert(pay, dsd_rate=2, out_rate=176400, bits=24, channels=2, stdout_mode=False, want_f64=False)
    assert out.read_bytes() == orc.WRITERS["W"](ref.planar, 24, 176400)
    # a filter without taps: the reference's message, non-zero exit, no file
    r = subprocess.run([cli, "-o", "W", "-r", "88200", "-t", "X", str(src)], capture_output=True, text=True)
    assert r.returncode == 1 and "Taps not found" in r.stderr and not (tmp_path / "pink.wav").exists()


@pytest.mark.gpu
@pytest.mark.parametrize("rel,flags,kw", [
    ("test/1kHz_stereo_i.dsd", ["-c", "2", "-f", "I", "-e", "M", "-r", "352800", "-b", "24", "-d", "X"],
     dict(channels=2, layout="I", lsb_first=False, dsd_rate=1, out_rate=352800, bits=24)),
    ("test/1kHz_mono_p.dsd", ["-c", "1", "-f", "P", "-e", "L", "-r", "96000", "-b", "32", "-d", "X"],
     dict(channels=1, layout="P", lsb_first=True, dsd_rate=1, out_rate=96000, bits=32)),
])
def test_cli_stdin_block_pushes_match_oracle(cli, ref_files, rel, flags, kw):
    """stdin -> stdout through d2p_push / d2p_pull in 4096-byte-per-channel blocks, like process_blocks."""
    raw = open(ref_files[rel], "rb").read()
    r = subprocess.run([cli] + flags + ["-"], input=raw, capture_output=True)
    assert r.returncode == 0, r.stderr.decode()
    ref = orclib.convert(np.frombuffer(raw, dtype=np.uint8), want_f64=False, **kw)
    assert r.stdout == ref.pcm.tobytes()
    assert b"seam:" in r.stderr and b"DSP speed for stdin" in r.stderr


@pytest.mark.gpu
def test_cli_output_directories_and_stdin_to_a_container(cli, ref_files, tmp_path):
    """derive_output_dir (rd/conversion_context.rs:447-477): with -p the input's directory, relative to the base directory
    (src/main.rs:262-273: the parent of the inputs' longest common path), is re-created under the output directory.
    stdin with a container output is written as output[_<rate>].<ext> (:418-425, convert_stdin src/main.rs:316-346)."""
    a, b = tmp_path / "lib" / "artist" / "one", tmp_path / "lib" / "artist" / "two"
    a.mkdir(parents=True); b.mkdir(parents=True)
    raw = open(ref_files["test/1kHz_stereo_128.dsf"], "rb").read()
    (a / "x.dsf").write_bytes(raw); (b / "y.dsf").write_bytes(raw)
    out = tmp_path / "out"
    r = subprocess.run([cli, "-o", "W", "-r", "176400", "-b", "24", "-d", "X", "-p", str(out), str(a / "x.dsf"), str(b / "y.dsf")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    # common path lib/artist, base directory lib: outputs land in out/artist/one and out/artist/two
    x, y = out / "artist" / "one" / "x.wav", out / "artist" / "two" / "y.wav"
    assert x.exists() and y.exists() and x.read_bytes() == y.read_bytes()
    # a single input: base directory = its own directory, so the file sits directly in -p
    out1 = tmp_path / "out1"
    r = subprocess.run([cli, "-o", "W", "-r", "176400", "-b", "24", "-d", "X", "-a", "-p", str(out1), str(a / "x.dsf")], capture_output=True, text=True)
    assert r.returncode == 0 and (out1 / "x_176_4K.wav").read_bytes() == x.read_bytes()
    # stdin -> WAV in -p, and in the working directory without -p
    dsd = open(ref_files["test/1kHz_stereo_i.dsd"], "rb").read()
    flags = ["-c", "2", "-f", "I", "-e", "M", "-r", "352800", "-b", "24", "-d", "X"]
    ref = orclib.convert(np.frombuffer(dsd, dtype=np.uint8), channels=2, layout="I", lsb_first=False, dsd_rate=1, out_rate=352800,
                         bits=24, stdout_mode=False, want_f64=False)
    r = subprocess.run([cli] + flags + ["-o", "W", "-a", "-p", str(out), "-"], input=dsd, capture_output=True)
    assert r.returncode == 0, r.stderr.decode()
    assert (out / "output_352_8K.wav").read_bytes() == orc.WRITERS["W"](ref.planar, 24, 352800)
    assert not list(out.glob("*.tmp"))
    cwd = tmp_path / "cwd"
    cwd.mkdir()
    r = subprocess.run([cli] + flags + ["-o", "A", "-"], input=dsd, capture_output=True, cwd=cwd)
    assert r.returncode == 0, r.stderr.decode()
    assert (cwd / "output.aif").read_bytes() == orc.WRITERS["A"](ref.planar, 24, 352800)


@pytest.mark.gpu
def test_cancel_flag_and_progress_callback(tmp_path):
    raw = np.random.default_rng(5).integers(0, 256, size=2 * 4096 * 60000, dtype=np.uint8)      # 8 chunks of 64 MiB
    cfg = d2p.make_config(dsd_rate=4, out_rate=352800, bits=24, channels=2)
    ref_frames = d2p.total_frames(cfg, raw.size)
    h_in, h_out = d2p.HostBuffer(raw.size), d2p.HostBuffer(ref_frames * 6)
    h_in.array[:] = raw
    batch = d2p.Batch(cfg)
    seen = []
    batch.set_control(None, lambda done, total: seen.append((done, total)))
    assert batch.convert_host(h_in.array, h_out.array) == ref_frames
    assert len(seen) >= 4 and seen[-1] == (ref_frames, ref_frames)
    assert all(a[0] <= b[0] for a, b in zip(seen, seen[1:])) and all(t == ref_frames for _, t in seen)
    whole = h_out.array.copy()
    # raise the flag from inside the progress callback: the call ends with D2P_ERR_CANCELLED a few chunks later
    flag = C.c_int32(0)
    calls = []

    def cancel_after_two(done, total):
        calls.append(done)
        if len(calls) == 2:
            flag.value = 1

    batch.set_control(flag, cancel_after_two)
    with pytest.raises(_capi.D2PError) as e:
        batch.convert_host(h_in.array, h_out.array)
    assert e.value.status == _capi.ERR_CANCELLED and len(calls) <= 3
    # the object stays usable, and an already-raised flag stops the call before any work
    flag.value = 0
    batch.set_control(flag, None)
    assert batch.convert_host(h_in.array, h_out.array) == ref_frames and np.array_equal(h_out.array, whole)
    batch.close()
    # file level: a cancelled conversion leaves no output behind
    src = tmp_path / "in.dsd"
    raw[: 2 * 4096 * 200].tofile(src)
    flag.value = 1
    with pytest.raises(_capi.D2PError) as e:
        d2p.convert_file(src, tmp_path / "o.wav", raw_channels=2, raw_dsd_rate=4, cancel_flag=flag)
    assert e.value.status == _capi.ERR_CANCELLED and not (tmp_path / "o.wav").exists()
    h_in.free(); h_out.free()


@pytest.mark.gpu
def test_per_job_peaks_for_level_batches():
    """dsd_levels prints one peak per file (src/bin/dsd_levels/main.rs:207): a device-resident batch reports per job."""
    import torch
    cfg = d2p.make_config(dsd_rate=1, out_rate=352800, bits=32, dither="X", out_format=_capi.OUT_NONE)
    raws = [np.random.default_rng(40 + i).integers(0, 256, size=2 * 4096 * (30 + 7 * i), dtype=np.uint8) for i in range(4)]
    raws[2][:] = 0x69                                                     # a silent file between loud ones
    batch = d2p.Batch(cfg)
    d_in = [torch.from_numpy(r).cuda() for r in raws]
    jobs = batch.make_jobs(dict(d_in=t.data_ptr(), in_bytes=t.numel(), d_out=None, out_cap_bytes=0) for t in d_in)
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(2):                                                    # the table describes the LAST run only
        batch.run(jobs, stream)
    peaks, clipped = batch.job_stats(len(raws), stream)
    refs = [orclib.check_level(r)[0] for r in raws]
    assert peaks == refs and clipped == [0, 0, 0, 0]
    assert batch.stats(stream)["peak_abs"] == max(refs)
    batch.stats_reset(stream)
    assert batch.stats(stream)["peak_abs"] == 0.0
    # the pipelined host conversion keeps no per-job table (its chunks are not the caller's jobs): asking says so
    batch.convert_host(raws[0], np.zeros(0, dtype=np.uint8))
    with pytest.raises(_capi.D2PError) as e:
        batch.job_stats(1, stream)
    assert e.value.status == _capi.ERR_INVALID_ARG
    assert batch.stats(stream)["peak_abs"] == refs[0]
    batch.close()


@pytest.mark.gpu
def test_convert_file_draws_a_seed_when_none_is_given(ref_files, tmp_path):
    """dither_seed 0 with a dither: seeded per file from the OS like thread_rng (rd/dither.rs:64-76); reported back."""
    src = ref_files["test/1kHz_stereo_128.dsf"]
    a = d2p.convert_file(src, tmp_path / "a.wav", out_rate=176400, bits=16, dither="T")
    b = d2p.convert_file(src, tmp_path / "b.wav", out_rate=176400, bits=16, dither="T")
    assert a["dither_seed"] != 0 and a["dither_seed"] != b["dither_seed"]
    assert (tmp_path / "a.wav").read_bytes() != (tmp_path / "b.wav").read_bytes()
    c = d2p.convert_file(src, tmp_path / "c.wav", out_rate=176400, bits=16, dither="T", dither_seed=a["dither_seed"])
    assert (tmp_path / "c.wav").read_bytes() == (tmp_path / "a.wav").read_bytes()          # reproducible from the report
    x = d2p.convert_file(src, tmp_path / "x.wav", out_rate=176400, bits=16, dither="X")
    assert x["dither_seed"] == 0


@pytest.mark.gpu
def test_convert_file_streams_long_files_in_bounded_slices(tmp_path):
    """A payload of several 64 MiB slices plus a ragged end goes through the streaming context with two fixed pinned
    buffers (the whole-payload staging needed 7 GB for a 60-minute multichannel file): the bytes equal the
    oracle's one-pass conversion and the progress callback walks up to the total."""
    per_ch = 4096 * 22000 + 1234                                   # 3 slices of 64 MiB and a partial block group
    raw = np.random.default_rng(9).integers(0, 256, size=2 * per_ch, dtype=np.uint8)
    src = tmp_path / "long.dsd"
    raw.tofile(src)
    seen = []
    res = d2p.convert_file(src, tmp_path / "o.wav", out_rate=176400, bits=24, container="W", raw_channels=2, raw_layout="P",
                           raw_lsb_first=True, raw_dsd_rate=4, progress=lambda done, total: seen.append((done, total)))
    ref = orclib.convert(raw, dsd_rate=4, out_rate=176400, bits=24, channels=2, stdout_mode=False, want_f64=False)
    assert res["frames"] == ref.frames[0]
    assert (tmp_path / "o.wav").read_bytes() == orc.WRITERS["W"](ref.planar, 24, 176400)
    assert len(seen) >= 3 and seen[-1] == (ref.frames[0], ref.frames[0]) and all(a[0] < b[0] for a, b in zip(seen, seen[1:]))
    # the reader / writer helpers that overlap neighbouring slices change nothing but the timing
    os.environ["D2P_FILE_SERIAL"] = "1"
    try:
        d2p.convert_file(src, tmp_path / "serial.wav", out_rate=176400, bits=24, container="W", raw_channels=2, raw_layout="P",
                         raw_lsb_first=True, raw_dsd_rate=4)
    finally:
        del os.environ["D2P_FILE_SERIAL"]
    assert (tmp_path / "serial.wav").read_bytes() == (tmp_path / "o.wav").read_bytes()
    assert res["peak_abs"] == np.abs(orclib.convert(raw, dsd_rate=4, out_rate=176400, bits=24, channels=2).f64 * (1 << 23)).max()
