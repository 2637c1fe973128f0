#!/usr/bin/env python
"""Benchmark of the UNAST spectral signal path on B200 (contract: task statement (4)).

    python bench.py --gpus N --steps K --warmup W             # our arm (CUDA path)
    python bench.py --impl reference --gpus N --steps K ...    # the reference's CPU algorithm (oracle port)

Primary workload (BASELINE.json configs[1], "C2"): the prepare_data.py-equivalent batch -- 13,100
synthetic LJSpeech-length utterances (1-10 s, PCM16 like LJSpeech) -> normalised mel + mag, sharded
over the N ranks longest-first with no collective (strong scaling: the job is fixed).  A "step" is
one pass of trim -> frame offsets -> fused front end over this rank's whole shard.
  value  audio-seconds (after trim) per second of device time, inputs and outputs resident in HBM
  e2e    the same job through unast_b200.pipeline.FrontendStream: pinned host PCM16 in, pinned host
         mel/mag out, H2D and D2H inside the timed region
Secondary workload ("gl", BASELINE.json configs[2], "C3"): spectrogram2wav on a batch of 64 magnitude
spectrograms, 100 Griffin-Lim iterations, power 1.2, de-emphasis + trim (one batch per rank: weak).
One JSON line is printed by rank 0.  oracle/ is used for the CPU baseline and a parity spot check only.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SR, HOP, NBINS, NMELS = 22050, 275, 1025, 80
FFT_FLOP = 56320.0                                  # 2.5 N log2 N, N = 2048
N_UTTS = 13100
METRIC = "audio-sec/s: mel front end & 100-iter Griffin-Lim at 1/2/4/8 B200; % of roofline"


# ---------------------------------------------------------------------------------------------
# synthetic workload
# ---------------------------------------------------------------------------------------------
def make_pool(n_pool, seed0=1000):
    from unast_b200 import synth
    return [synth.to_pcm16(synth.speech_like(10.0, seed0 + i)) for i in range(n_pool)]


def utterance(pool, i, n):
    """PCM16 utterance i of n samples: a prefix of a pool clip, per-utterance gain, quiet lead/tail kept."""
    base = pool[i % len(pool)]
    lead = min(int(0.25 * SR), n // 8)
    y = base[:n].astype(np.float32)
    y[n - lead:] = base[base.size - lead:]
    gain = 0.5 + 0.5 * ((i * 2654435761) % 1000) / 999.0
    return np.round(y * gain).astype(np.int16)


def c2_lengths():
    from unast_b200 import synth
    return np.round(synth.ljspeech_durations(N_UTTS, seed=1) * SR).astype(np.int64)


# ---------------------------------------------------------------------------------------------
# CPU arm (the reference's algorithm = oracle port); runs in forked workers, one BLAS thread each
# ---------------------------------------------------------------------------------------------
_POOL_CLIPS = None


def _cpu_init():
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        pass


def _cpu_frontend(args):
    i, n = args
    from oracle import spectral_oracle as so
    y = utterance(_POOL_CLIPS, i, n).astype(np.float32) / np.float32(32768.0)
    mel, mag, b = so.spectrograms_from_wave(y, return_bounds=True)
    return (b[1] - b[0]) / SR


def _cpu_gl(args):
    i, n, iters = args
    from oracle import spectral_oracle as so
    y = utterance(_POOL_CLIPS, i, n).astype(np.float32) / np.float32(32768.0)
    mag = so.spectrograms_from_wave(y)[1]
    t0 = time.perf_counter()
    so.spectrogram2wav(mag, iters=iters)
    return HOP * (mag.shape[0] - 1) / SR, time.perf_counter() - t0


class CpuArm:
    def __init__(self, cores=None):
        import multiprocessing as mp
        global _POOL_CLIPS
        self.cores = cores or len(os.sched_getaffinity(0)) or 1
        _POOL_CLIPS = make_pool(4)
        self.pool = mp.get_context("fork").Pool(self.cores, initializer=_cpu_init)
        self.lengths = c2_lengths()

    def frontend_sample(self, n_clips):
        idx = list(range(0, N_UTTS, max(1, N_UTTS // n_clips)))[:n_clips]
        jobs = [(i, int(self.lengths[i])) for i in idx]
        t0 = time.perf_counter()
        secs = sum(self.pool.map(_cpu_frontend, jobs, chunksize=1))
        return secs, time.perf_counter() - t0, len(jobs)

    def gl_sample(self, n_clips, iters, dur_s=1.5):
        jobs = [(i, int(dur_s * SR), iters) for i in range(n_clips)]
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_gl, jobs, chunksize=1)
        wall = time.perf_counter() - t0
        return sum(r[0] for r in res), wall, len(jobs)

    def close(self):
        self.pool.close()
        self.pool.join()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm()
    n_clips = max(32, 16 * arm.cores) if args.sample_clips is None else args.sample_clips
    for _ in range(args.warmup):
        arm.frontend_sample(max(arm.cores, n_clips // 8))
    secs = wall = 0.0
    for _ in range(args.steps):
        s, w, _ = arm.frontend_sample(n_clips)
        secs += s
        wall += w
    value = secs / wall
    gl_secs, gl_wall, gl_n = arm.gl_sample(arm.cores, args.gl_iters)
    arm.close()
    sample = "%d of the 13,100 C2 utterances per step (every %d-th), oracle port in %d forked workers, 1 BLAS thread each" % (
        n_clips, N_UTTS // n_clips, arm.cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C2 front end: 13,100 synthetic LJSpeech-length utterances -> mel+mag (bounded sample per step)",
                   "sample": sample},
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": arm.cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gl": {"value": gl_secs / gl_wall, "unit": "audio-s/s", "cores": arm.cores, "kind": "port",
               "sample": "%d clips of 1.5 s, %d classic Griffin-Lim iterations, one clip per worker" % (gl_n, args.gl_iters)},
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# clocks (NVML, sampled in a thread during the timed regions)
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting",
               0x10: "sync_boost"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        self._stop = threading.Event()
        self._active = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:                                   # report, never fake
            self.err = repr(e)
        self.th = threading.Thread(target=self._loop, daemon=True)
        if self.ok:
            self.th.start()

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            if self._active.is_set():
                try:
                    self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                    except Exception:
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in self.REASONS.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
                time.sleep(0.004)
            else:
                time.sleep(0.001)

    def __enter__(self):
        self._active.set()
        return self

    def __exit__(self, *a):
        self._active.clear()

    def summary(self):
        self._stop.set()
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "err", "?")]}
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return json.load(fh), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def ncu_traffic(precision, frames):
    """DRAM bytes (read + write) of one front-end launch from the committed `ncu --set full` capture
    (profiles/roofline_traffic.json), scaled per frame to this launch; None if no capture is recorded."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(p) as fh:
            rec = json.load(fh)["frontend_" + precision]
        return rec["dram_bytes"] / rec["frames"] * frames
    except Exception:
        return None


def gl_traffic(frames):
    """DRAM bytes of one Griffin-Lim iteration launch from the committed ncu capture, scaled per frame; None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as fh:
            rec = json.load(fh)["gl_iteration"]
        return rec["dram_bytes"] / rec["frames"] * frames
    except Exception:
        return None


def run_ours(args):
    import torch
    from unast_b200 import dist as D
    from unast_b200 import sharding
    rank, local_rank, world = D.env_world()
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = CpuArm()                                            # fork the CPU workers before CUDA exists
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    D.init_from_env("nccl")
    from unast_b200.engine import get_engine
    from unast_b200.pipeline import FrontendStream, plan_chunks
    eng = get_engine(local_rank)
    dev = eng.device
    peaks, peak_src = measured_peaks()
    clocks = ClockSampler(local_rank)

    # ---- workload C2: this rank's shard of the 13,100 utterances --------------------------------
    lengths_all = c2_lengths()
    if args.utts:
        lengths_all = lengths_all[:args.utts]
    shard = sharding.deal_longest_first(1 + lengths_all // HOP, world)[rank]
    shard = sorted(shard, key=lambda i: (-int(lengths_all[i]), i))        # length-bucketed order
    lengths = lengths_all[shard]
    pool = make_pool(args.pool)
    offsets_np = np.concatenate([[0], np.cumsum(lengths)])
    wave_host = torch.empty(int(offsets_np[-1]), dtype=torch.int16).pin_memory()
    wh = wave_host.numpy()
    for j, i in enumerate(shard):
        wh[offsets_np[j]:offsets_np[j + 1]] = utterance(pool, i, int(lengths[j]))
    B = len(shard)
    max_len = int(lengths.max())
    frames_ub = int((1 + lengths // HOP).sum())

    wave = wave_host.to(dev, non_blocking=True)
    offsets = torch.from_numpy(offsets_np).to(dev)
    mel = torch.empty(NMELS * frames_ub, dtype=torch.float32, device=dev)
    mag = torch.empty((frames_ub, NBINS), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()

    fe_ms = []

    def step():
        bounds = eng.trim_bounds(wave, offsets, max_len)
        fo = eng.frame_offsets(bounds)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.frontend(wave, offsets, bounds, fo, 0, precision=args.precision, mel_out=mel, mag_out=mag)
        e1.record()
        fe_ms.append((e0, e1))
        return bounds, fo

    for _ in range(args.warmup):
        bounds, fo = step()
    eng.check()
    fo_h = fo.cpu().numpy()
    bd_h = bounds.cpu().numpy()
    my_frames = int(fo_h[-1])
    my_audio_s = float((bd_h[:, 1] - bd_h[:, 0]).sum()) / SR
    fe_ms.clear()
    launches0 = eng.launches
    D.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with clocks:
        t0.record()
        for _ in range(args.steps):
            step()
        t1.record()
        torch.cuda.synchronize()
    D.barrier()
    step_ms = D.max_over_ranks(t0.elapsed_time(t1)) / args.steps
    launches = eng.launches - launches0
    fe_kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in fe_ms]))
    total_audio_s = D.sum_over_ranks(my_audio_s)
    total_frames = D.sum_over_ranks(my_frames)
    value = total_audio_s / (step_ms * 1e-3)

    # roofline of the dominant kernel (fused front end): algorithmic bytes / frame = 550 (PCM16 in) + 4100 + 320
    bytes_per_frame = 275 * 2 + NBINS * 4 + NMELS * 4
    achieved = my_frames * bytes_per_frame / (fe_kernel_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "traffic": ncu_traffic(args.precision, my_frames),
                "kernel": "frontend_au_kernel<double,12>" if args.precision == "f64" else "frontend_au_kernel<float,16>",
                "kernel_ms": fe_kernel_ms, "frames_per_launch": my_frames, "bytes_per_frame": bytes_per_frame,
                "peak_source": peak_src,
                "fft_tflops": my_frames * FFT_FLOP / (fe_kernel_ms * 1e-3) / 1e12}

    # parity spot check of the timed buffers against the oracle (outside the timed region)
    spot = None
    if rank == 0:
        from oracle import spectral_oracle as so
        spot = {"utterances": [], "bounds_equal": True, "max_abs_mag": 0.0, "max_abs_mel": 0.0}
        for j in sorted({0, B // 4, B // 2, B - 1}):                          # first, quarter, middle, last of the shard
            y = wh[offsets_np[j]:offsets_np[j + 1]].astype(np.float32) / np.float32(32768.0)
            omel, omag, ob = so.spectrograms_from_wave(y, return_bounds=True)
            T = int(fo_h[j + 1] - fo_h[j])
            gmag = mag[int(fo_h[j]):int(fo_h[j + 1])].cpu().numpy()
            gmel = mel[NMELS * int(fo_h[j]):NMELS * int(fo_h[j + 1])].cpu().numpy().reshape(NMELS, T).T
            spot["utterances"].append(int(shard[j]))
            spot["bounds_equal"] = spot["bounds_equal"] and bool(tuple(int(v) for v in bd_h[j]) == ob)
            if gmag.shape != omag.shape or gmel.shape != omel.shape:
                spot["max_abs_mag"] = spot["max_abs_mel"] = None
                break
            spot["max_abs_mag"] = max(spot["max_abs_mag"], float(np.abs(gmag - omag).max()))
            spot["max_abs_mel"] = max(spot["max_abs_mel"], float(np.abs(gmel - omel).max()))

    # ---- e2e: pinned host in -> pinned host out through the public streaming API ---------------
    del mel, mag
    torch.cuda.empty_cache()
    chunks = plan_chunks(lengths, max_utts=args.chunk_utts, max_samples=args.chunk_utts * 150000)
    stream = FrontendStream(lengths, chunks=chunks, precision=args.precision)
    stream.run(wave_host)                                         # warm-up pass (allocations, page-in)
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    stream.h2d_bytes = stream.d2h_bytes = 0
    D.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(e2e_steps):
        stream.run(wave_host)
    for s in (stream.s_in, stream.s_comp, stream.s_out):
        torch.cuda.current_stream().wait_stream(s)
    t1.record()
    torch.cuda.synchronize()
    D.barrier()
    e2e_ms = D.max_over_ranks(t0.elapsed_time(t1)) / e2e_steps
    e2e = {"value": total_audio_s / (e2e_ms * 1e-3), "unit": "audio-s/s",
           "h2d_bytes_per_step": int(D.sum_over_ranks(stream.h2d_bytes / e2e_steps)),
           "d2h_bytes_per_step": int(D.sum_over_ranks(stream.d2h_bytes / e2e_steps)),
           "ms_per_step": e2e_ms, "steps": e2e_steps,
           "api": "unast_b200.pipeline.FrontendStream.run (pinned PCM16 in, pinned mel/mag out, %d chunks)" % len(chunks)}
    del stream
    torch.cuda.empty_cache()

    # ---- secondary workload C3: 100-iteration Griffin-Lim on 64 spectrograms per rank -----------
    gl = run_gl(args, eng, clocks, pool, rank, world, D)
    loader = run_loader(args, eng, pool, rank) if rank == 0 else None
    longform = run_longform(args, eng, pool, rank, world, D)

    # ---- pipe peaks (roofline denominators MEASURED_PEAKS.json lacks) ---------------------------
    pipes = {k: eng.pipe_peak_tflops(k) for k in ("f32", "f32x2", "f64")}
    if gl is not None:
        peak32 = max(pipes["f32"], pipes["f32x2"])
        gl["roofline"]["peak"] = peak32
        gl["roofline"]["frac"] = gl["roofline"]["achieved"] / peak32
        if gl.get("fast"):
            gl["fast"]["roofline_frac"] = gl["fast"]["roofline_frac"] / peak32
    roofline["fp64_peak_tflops"] = pipes["f64"]
    if args.precision == "f64":
        roofline["frac_of_fp64_pipe"] = roofline["fft_tflops"] / pipes["f64"]

    cpu_baseline = None
    if cpu is not None:
        n_clips = max(32, 16 * cpu.cores)
        cpu.frontend_sample(cpu.cores)                            # warm the workers
        secs, wall, n = cpu.frontend_sample(n_clips)
        gsecs, gwall, gn = cpu.gl_sample(cpu.cores, args.gl_iters)
        cpu.close()
        cpu_baseline = {"value": secs / wall, "unit": "audio-s/s", "cores": cpu.cores, "kind": "port",
                        "sample": "%d of the 13,100 C2 utterances (every %d-th), oracle port (numpy.fft float64, BLAS sgemm) in %d forked "
                                  "workers, 1 BLAS thread each" % (n, N_UTTS // n_clips, cpu.cores),
                        "gl": {"value": gsecs / gwall, "unit": "audio-s/s",
                               "sample": "%d clips of 1.5 s, %d classic Griffin-Lim iterations, one per worker" % (gn, args.gl_iters)}}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "C2 (BASELINE.json configs[1]): %d synthetic LJSpeech-length utterances (1-10 s, PCM16) -> "
                                   "normalised mel+mag, sharded longest-first over %d rank(s), no collective" % (len(lengths_all), world),
                       "utterances": int(len(lengths_all)), "frames": int(total_frames), "audio_s_after_trim": total_audio_s,
                       "l2": "inputs+outputs per step are %.1f GB >> 126 MB L2 (no flush needed)" % (
                           (my_frames * bytes_per_frame) / 1e9),
                       "fft_precision": args.precision, "pool_clips": args.pool},
            "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
            "cpu_baseline": cpu_baseline, "gl": gl, "loader": loader, "longform": longform, "pipe_peaks_tflops": pipes, "parity_spot_check": spot,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def run_longform(args, eng, pool, rank, world, D):
    """C5 (BASELINE.json configs[4]): one long waveform; the front end's frames are sharded over the ranks with a halo
    (unast_b200.longform, no exchange); the 100-iteration Griffin-Lim round trip of the whole clip runs on rank 0 only
    (replicas: sharding it would need a seam exchange per iteration).  Latency regime -- reported, not headlined."""
    import torch
    from unast_b200 import longform as LF
    secs = args.longform_seconds
    if secs <= 0:
        return None
    n = int(secs * SR)
    reps = -(-n // pool[0].size)
    w = np.concatenate([pool[(3 * i) % len(pool)] for i in range(reps)])[:n].astype(np.int16)
    T = 1 + n // HOP
    a, b = LF.frame_ranges(T, world)[rank]
    c0, c1, keep = LF.chunk_for_range(n, a, b)
    dev = eng.device
    chunk = torch.from_numpy(w[c0:c1].copy()).to(dev)
    nc = c1 - c0
    off = torch.tensor([0, nc], dtype=torch.int64, device=dev)
    bounds = torch.tensor([[0, nc]], dtype=torch.int32, device=dev)
    fo = eng.frame_offsets(bounds)
    Tc = int(fo.cpu().numpy()[-1])
    mel = torch.empty(NMELS * Tc, dtype=torch.float32, device=dev)
    mag = torch.empty((Tc, NBINS), dtype=torch.float32, device=dev)
    for _ in range(3):
        eng.frontend(chunk, off, bounds, fo, Tc, precision=args.precision, mel_out=mel, mag_out=mag)
    D.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(10):
        eng.frontend(chunk, off, bounds, fo, Tc, precision=args.precision, mel_out=mel, mag_out=mag)
    t1.record()
    torch.cuda.synchronize()
    D.barrier()
    fe_ms = D.max_over_ranks(t0.elapsed_time(t1) / 10)
    out = {"workload": "C5 (BASELINE.json configs[4]): %d s waveform, %d frames; front end sharded by frame range over %d rank(s) "
                       "with a %d-frame halo, no exchange" % (secs, T, world, LF.HALO),
           "frontend": {"value": secs / (fe_ms * 1e-3), "unit": "audio-s/s", "ms": fe_ms, "frames_this_rank": int(b - a)}}
    if rank == 0 and world == 1 and args.gl_iters > 0:
        frames = np.array([Tc], dtype=np.int32)
        wav_out = torch.empty(HOP * (Tc - 1), dtype=torch.float32, device=dev)
        step = lambda: eng.griffinlim(mag, frames, n_iter=args.gl_iters, momentum=0.0, power=1.2, deemphasis=True, want_trim=True,
                                      wav_out=wav_out)
        step()
        torch.cuda.synchronize()
        t0.record()
        step()
        t1.record()
        torch.cuda.synchronize()
        gl_ms = t0.elapsed_time(t1)
        out["gl_round_trip"] = {"value": secs / (gl_ms * 1e-3), "unit": "audio-s/s", "ms": gl_ms, "iters": args.gl_iters,
                                "multi_gpu": "replicas only"}
    return out


def run_loader(args, eng, pool, rank):
    """C4 (BASELINE.json configs[3]): on-the-fly mel front end inside the loader -- collate_fn_transformer semantics on a
    batch of 32 utterances (sorted by text length, zero padded [32, T_max, 80] device tensor), through the public
    unast_b200.preprocess API with host PCM16 in (H2D inside the timed region) and the mel lengths read back."""
    import torch
    from unast_b200 import preprocess as PP, synth
    if args.loader_batch <= 0:
        return None
    B = args.loader_batch
    rng = np.random.RandomState(4)
    lens = np.round(synth.ljspeech_durations(B, seed=4) * SR).astype(np.int64)
    batch = [{"text": rng.randint(1, 70, size=int(rng.randint(20, 150))).astype(np.int32), "wave": utterance(pool, 11 * j + 3, int(n))}
             for j, n in enumerate(lens)]
    for d in batch:
        d["text_length"] = len(d["text"])
    for _ in range(3):
        out = PP.collate_fn_transformer(batch)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    t0.record()
    for _ in range(n):
        text, mel, text_length, mel_length = PP.collate_fn_transformer(batch)
        ml = mel_length.cpu()
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / n
    audio_s = float(HOP * (ml.numpy() - 1).sum()) / SR
    # next row (8(f) rank 2): specaugment + noise_fn on that batch (train.py:236, network.py:205), and the masking
    # kernel against HBM on a batch large enough to leave the latency regime (its roofline: read + write 4 B/element,
    # plus one more read for the slab mean)
    from unast_b200 import augment as AUG
    for _ in range(3):
        aug = AUG.specaugment_noise(mel, ml)
    torch.cuda.synchronize()
    t0.record()
    for _ in range(n):
        aug = AUG.specaugment_noise(mel, ml)
    t1.record()
    torch.cuda.synchronize()
    aug_ms = t0.elapsed_time(t1) / n
    Bb, Tb = 1024, 800
    big = torch.rand((Bb, Tb, 80), device=mel.device)
    big_out = torch.empty_like(big)
    spans = torch.from_numpy(np.tile(np.array([[5, 19, 100, 99]], np.int32), (Bb, 1))).to(mel.device)
    keep = (torch.rand((Bb, Tb), device=mel.device) > 0.3).float()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=mel.device)
    tot = 0.0
    for r in range(6):
        flush.zero_()
        t0.record()
        eng.mel_augment(big, spans=spans, keep=keep, out=big_out)
        t1.record()
        torch.cuda.synchronize()
        if r:
            tot += t0.elapsed_time(t1)
    big_ms = tot / 5
    alg = 12.0 * big.numel()
    peak = measured_peaks()[0]["hbm_gbs"]
    del big, big_out, flush
    return {"workload": "C4 (BASELINE.json configs[3]): collate_fn_transformer with on-the-fly mels, batch %d, PCM16 host -> "
                        "[%d, %d, 80] float32 device tensor" % (B, B, int(mel.shape[1])),
            "e2e": {"value": audio_s / (ms * 1e-3), "unit": "audio-s/s", "ms_per_batch": ms,
                    "h2d_bytes_per_step": int(2 * lens.sum()), "d2h_bytes_per_step": int(8 * B + 4 * (B + 1))},
            "augment": {"workload": "specaugment + noise_fn in one pass (unast_mel_augment), host draws included",
                        "ms_per_batch": aug_ms,
                        "kernel": {"workload": "[%d, %d, 80] float32, spans + keep mask, L2 flushed between launches" % (Bb, Tb),
                                   "ms": big_ms, "algorithmic_bytes": alg, "achieved_gbps": alg / (big_ms * 1e-3) / 1e9,
                                   "peak_gbps": peak, "frac": (alg / (big_ms * 1e-3) / 1e9 / peak) if peak else None}},
            "note": "latency regime: one 32-utterance batch is ~15 k frames, a tenth of what fills the GPU"}


def run_gl(args, eng, clocks, pool, rank, world, D):
    """C3: utils.spectrogram2wav semantics (classic GL from zero phase) on 64 spectrograms, n_iter = 100."""
    import torch
    from unast_b200 import synth
    if args.gl_batch <= 0:
        return None
    dev = eng.device
    durs = synth.ljspeech_durations(args.gl_batch, seed=2)
    lens = np.round(durs * SR).astype(np.int64)
    waves = [utterance(pool, 7 * j + rank, int(n)) for j, n in enumerate(lens)]
    off_np = np.concatenate([[0], np.cumsum(lens)])
    wave = torch.from_numpy(np.concatenate(waves)).to(dev)
    off = torch.from_numpy(off_np).to(dev)
    bounds = eng.trim_bounds(wave, off, int(lens.max()))
    fo = eng.frame_offsets(bounds)
    eng.check()
    fo_h = fo.cpu().numpy()
    _, mag = eng.frontend(wave, off, bounds, fo, int(fo_h[-1]), precision="f64")
    frames = np.diff(fo_h).astype(np.int32)
    audio_s = float(HOP * (frames - 1).sum()) / SR
    n_frames = int(frames.sum())
    wav_out = torch.empty(int(HOP * (frames.astype(np.int64) - 1).sum()), dtype=torch.float32, device=dev)

    def step():
        return eng.griffinlim(mag, frames, n_iter=args.gl_iters, momentum=0.0, power=1.2, deemphasis=True, want_trim=True,
                              wav_out=wav_out)
    for _ in range(max(1, args.warmup // 2)):
        step()
    eng.check()
    gl_steps = max(1, min(args.steps, args.gl_steps))
    l0 = eng.launches
    D.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with clocks:
        t0.record()
        for _ in range(gl_steps):
            step()
        t1.record()
        torch.cuda.synchronize()
    D.barrier()
    ms = D.max_over_ranks(t0.elapsed_time(t1)) / gl_steps
    launches = (eng.launches - l0) // gl_steps
    total_audio = D.sum_over_ranks(audio_s)
    flop = (2 * args.gl_iters + 1) * FFT_FLOP * n_frames
    # e2e: host magnitudes in (pinned), host waveforms out
    mag_h = mag.cpu().pin_memory()
    wav_h = torch.empty(wav_out.numel(), dtype=torch.float32).pin_memory()
    bd_h = torch.empty((len(frames), 2), dtype=torch.int32).pin_memory()
    mag_d = torch.empty_like(mag)

    def e2e_step():
        mag_d.copy_(mag_h, non_blocking=True)
        w, _, b = eng.griffinlim(mag_d, frames, n_iter=args.gl_iters, momentum=0.0, power=1.2, deemphasis=True, want_trim=True,
                                 wav_out=wav_out)
        wav_h.copy_(w, non_blocking=True)
        bd_h.copy_(b, non_blocking=True)
    e2e_step()
    D.barrier()
    torch.cuda.synchronize()
    t0.record()
    for _ in range(gl_steps):
        e2e_step()
    t1.record()
    torch.cuda.synchronize()
    D.barrier()
    e2e_ms = D.max_over_ranks(t0.elapsed_time(t1)) / gl_steps
    # the production variant of src/gl_vocoder.py:26 on the same batch: librosa.griffinlim, 32 iterations, momentum 0.99,
    # seeded random initial phases (device-timed only)
    fast = None
    if args.gl_iters > 0:
        rs = np.random.RandomState(3)
        ang = np.exp(2j * np.pi * rs.rand(n_frames, NBINS)).astype(np.complex64)
        a0 = torch.from_numpy(ang).contiguous().to(dev)

        def fast_step():
            return eng.griffinlim(mag, frames, n_iter=32, momentum=0.99, angles0=a0, power=1.2, deemphasis=True, want_trim=True,
                                  wav_out=wav_out)
        fast_step()
        torch.cuda.synchronize()
        t0.record()
        for _ in range(gl_steps):
            fast_step()
        t1.record()
        torch.cuda.synchronize()
        fms = D.max_over_ranks(t0.elapsed_time(t1)) / gl_steps
        fast = {"workload": "gl_vocoder.py variant: 32 iterations, momentum 0.99, seeded phases, same batch", "ms_per_step": fms,
                "value": total_audio / (fms * 1e-3), "unit": "audio-s/s",
                "roofline_frac": (2 * 32 + 1) * FFT_FLOP * n_frames / (fms * 1e-3) / 1e12}
    return {"fast": fast,
            "workload": "C3 (BASELINE.json configs[2]): spectrogram2wav, %d classic Griffin-Lim iterations, power 1.2, batch of %d "
                        "spectrograms per rank (weak scaling), de-emphasis + trim" % (args.gl_iters, args.gl_batch),
            "value": total_audio / (ms * 1e-3), "unit": "audio-s/s", "ms_per_step": ms, "steps": gl_steps, "scaling": "weak",
            "frames_per_rank": n_frames, "gpu_launches_per_step": int(launches), "dtype": "f32",
            "roofline": {"bound": "fp32", "achieved": flop / (ms * 1e-3) / 1e12, "peak": None, "unit": "TFLOP/s", "frac": None,
                         "kernel": "gl_pk_kernel<false>", "traffic": gl_traffic(n_frames),
                         "flop_per_frame": (2 * args.gl_iters + 1) * FFT_FLOP,
                         "note": "whole step (all %d launches); algorithmic flop = (2 n_iter + 1) x 2.5 N log2 N per frame" % launches},
            "e2e": {"value": total_audio / (e2e_ms * 1e-3), "unit": "audio-s/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(mag_h.numel() * 4 * world), "d2h_bytes_per_step": int((wav_h.numel() * 4 + bd_h.numel() * 4) * world)}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--precision", choices=["f64", "f32"], default="f64",
                    help="front-end FFT arithmetic; f64 is the parity mode (reference: numpy.fft float64)")
    ap.add_argument("--utts", type=int, default=0, help="use only the first N of the 13,100 utterances (debug)")
    ap.add_argument("--pool", type=int, default=24, help="distinct 10 s synthetic clips the utterances are cut from")
    ap.add_argument("--chunk-utts", type=int, default=512)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--gl-batch", type=int, default=64)
    ap.add_argument("--gl-iters", type=int, default=100)
    ap.add_argument("--gl-steps", type=int, default=3)
    ap.add_argument("--loader-batch", type=int, default=32)
    ap.add_argument("--longform-seconds", type=int, default=600)
    ap.add_argument("--sample-clips", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        print("bench.py: warning: fewer than 3 warm-up steps", file=sys.stderr)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
