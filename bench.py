#!/usr/bin/env python
"""Headline benchmark: audio-seconds per second (RTFx) of log-mel + Whisper audio-encoder.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]

One "step" = one pass of the hot path (PCM -> log-mel -> normalise -> encoder states) over one batch of
synthetic 30 s clips.  Default workload: BASELINE.json configs[2]'s per-GPU shard -- Whisper-small,
256 clips per GPU, bf16 tensor-core compute -- the configuration the north_star target
(>= 100 000x realtime on 8 GPUs) is quoted on; per-GPU work is fixed as N grows (weak scaling, clips
sharded by rank, no collective on the data path).  Other BASELINE configs: --workload base64 | tiny1 |
largev3 | frontend.

The JSON line (rank 0) carries: `value` (PCM already resident in HBM, CUDA events, max over ranks),
`e2e` (pinned host PCM -> H2D -> kernels -> D2H of the encoder states inside the timed region, through
yoho_b200.pipeline.ClipEncoder.encode_host), `roofline` for the dominant kernel class (tcgen05 GEMM)
from per-launch CUDA events over the same timed region, `rooflines` for the other classes, `clocks`
sampled by nvidia-smi during the timed region, `gpu_launches`, and `cpu_baseline` (the oracle timed on
the host cores; N=1 only).  `--impl reference` times that CPU oracle alone (the reference's own JAX code
cannot run on this image: jax/flax are absent -- DESIGN.md).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

CLIP_SECONDS = 30.0
N_SAMPLES = 480000
GFLOP_PER_CLIP = {"tiny": 36.94, "base": 87.37, "small": 344.16, "large-v3": 2273.77}   # SURVEY 8a / DESIGN.md

WORKLOADS = {
    # name: (model, clips per GPU, stages, description)
    "small256": ("small", 256, "logmel+encoder", "BASELINE configs[2] per-GPU shard: whisper-small, 256 x 30 s clips per GPU"),
    "base64": ("base", 64, "logmel+encoder", "BASELINE configs[1]: whisper-base, 64 x 30 s clips"),
    "tiny1": ("tiny", 1, "logmel+encoder", "BASELINE configs[0]: whisper-tiny, one 30 s clip"),
    "largev3": ("large-v3", 64, "encoder", "BASELINE configs[3]: whisper-large-v3 encoder only, 64 clips per GPU"),
    "frontend": ("small", 960, "logmel", "BASELINE configs[4]: log-mel only, 8 h of PCM (960 clips) per GPU"),
}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_burst": float(d["bf16_tflops"]),
                "bf16_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi sampler running beside the timed region (B200_PROFILING.md clocks line)."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,timestamp")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.t_mark = gpu_index, None, None

    def mark(self):
        """Only samples taken from now on count (the process is started during the warm-up: nvidia-smi's start-up is ~1 s of
        driver activity that measurably slows the kernels it runs beside)."""
        import datetime
        self.t_mark = datetime.datetime.now()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        return self.summarise(out, self.t_mark)

    @staticmethod
    def summarise(out: str, t_mark=None) -> dict:
        """nvidia-smi csv lines -> the `clocks` object; only lines stamped at or after t_mark count (all, if none is)."""
        import datetime
        lines = [[x.strip() for x in line.split(",")] for line in out.splitlines()]
        lines = [f for f in lines if len(f) >= 9]

        def inside(f):
            if t_mark is None or len(f) < 10:
                return True
            try:
                return datetime.datetime.strptime(f[9], "%Y/%m/%d %H:%M:%S.%f") >= t_mark
            except ValueError:
                return True
        timed_lines = [f for f in lines if inside(f)]
        window = "timed region"
        if not timed_lines and lines:            # a timed region shorter than the sampling period: the warm-up just before it
            timed_lines, window = lines[-5:], "end of warm-up + timed region (timed region shorter than the 0.2 s sampling period)"
        sm, smax, power, reasons = [], [], [], set()
        for f in timed_lines:
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s, p in zip(sm, power) if p >= 0.5 * max(power)] or sm
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": max(smax), "power_w_max": max(power),
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------ CPU oracle timing
def cpu_oracle_rtfx(model: str, stages: str, target_seconds: float = 12.0, max_clips: int = 16):
    """Times the CPU oracle (oracle/: scipy/numpy frontend restatement + torch-CPU fp32 encoder restatement,
    all host threads) on a bounded sample of the workload.  Returns (rtfx, cores, sample description)."""
    import numpy as np
    import torch
    from oracle import encoder as eo
    from oracle import frontend as fo
    from yoho_b200 import MODEL_SHAPES
    from yoho_b200.synth import synth_batch

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_mels, dims, heads, layers = MODEL_SHAPES[model]
    params = eo.init_params(n_mels, dims, layers, seed=42) if "encoder" in stages else None
    pos = eo.sinusoid_table(1500, dims) if params is not None else None

    def run(pcm):
        if "logmel" in stages:
            mel = fo.log_mel(pcm, n_mels=n_mels, norm_mode=fo.NORM_REFERENCE, dtype=np.float32)
        else:
            mel = np.zeros((pcm.shape[0], 2999, n_mels), np.float32) + pcm[:, :1, None]
        if params is not None:
            return eo.audio_encoder_torch_cpu(params, mel, heads, pos)
        return mel

    one = synth_batch(1, N_SAMPLES)
    run(one)                                             # warm-up (thread pools, FFT plans)
    t0 = time.perf_counter(); run(one); t1 = time.perf_counter() - t0
    n = int(max(1, min(max_clips, round(target_seconds / max(t1, 1e-3)))))
    batch = min(n, 4)
    pcm = synth_batch(batch, N_SAMPLES, unique=min(batch, 4))
    t0 = time.perf_counter()
    done = 0
    while done < n:
        run(pcm)
        done += batch
    dt = time.perf_counter() - t0
    sample = f"{done} clips ({done * CLIP_SECONDS:.0f} s of audio) in batches of {batch}, {model} {stages}, fp32, {dt:.1f} s wall"
    return done * CLIP_SECONDS / dt, cores, sample, dt / done


def run_reference_arm(args):
    """`--impl reference`: the CPU restatement of the reference path on the box's host cores."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return 0
    model, clips, stages, desc = WORKLOADS[args.workload]
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    # one step = a bounded sample: clips_per_step clips through the oracle; size it from one timed clip
    rtfx0, cores, _, sec_per_clip = cpu_oracle_rtfx(model, stages, target_seconds=4.0, max_clips=4)
    budget = 150.0 / max(1, args.steps + args.warmup)
    per_step = int(max(1, min(8, budget / max(sec_per_clip, 1e-3))))
    import numpy as np
    from oracle import encoder as eo
    from oracle import frontend as fo
    from yoho_b200 import MODEL_SHAPES
    from yoho_b200.synth import synth_batch
    n_mels, dims, heads, layers = MODEL_SHAPES[model]
    params = eo.init_params(n_mels, dims, layers, seed=42) if "encoder" in stages else None
    pos = eo.sinusoid_table(1500, dims) if params is not None else None
    pcm = synth_batch(per_step, N_SAMPLES, unique=min(per_step, 4))

    def step():
        mel = fo.log_mel(pcm, n_mels=n_mels, norm_mode=fo.NORM_REFERENCE, dtype=np.float32) if "logmel" in stages \
            else np.zeros((per_step, 2999, n_mels), np.float32)
        return eo.audio_encoder_torch_cpu(params, mel, heads, pos) if params is not None else mel

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = per_step * args.steps * CLIP_SECONDS / dt
    sample = f"{per_step} clips per step ({model}, {stages}, fp32 oracle: numpy/scipy frontend + torch-CPU encoder), {args.steps} steps"
    line = {"impl": "reference", "metric": "audio-sec/sec (RTFx) log-mel+encoder", "value": value, "unit": "x realtime",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "clips_per_step": per_step, "note": "reference JAX code not runnable here (jax/flax absent); CPU oracle port timed instead"},
            "cpu_baseline": {"value": value, "unit": "x realtime", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "x realtime", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps (default 10; 2000 for the single-clip tiny1 workload so that the timed region is long enough to sample clocks)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="small256", choices=sorted(WORKLOADS))
    ap.add_argument("--clips", type=int, default=0, help="override clips per GPU")
    ap.add_argument("--chunk", type=int, default=0, help="clips per kernel-sequence chunk on the device path (0 = whole batch)")
    ap.add_argument("--e2e-chunk", type=int, default=64, help="clips per pipelined chunk on the host path")
    ap.add_argument("--norm", default="reference", choices=["reference", "whisper"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="tiny1: plain launches instead of the CUDA-graph replay")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3                                   # timing rule: W >= 3
    if args.steps <= 0:
        args.steps = 2000 if (args.workload == "tiny1" and args.impl == "ours") else 10
    if args.impl == "reference":
        return run_reference_arm(args)

    import numpy as np
    import torch
    from yoho_b200 import MODEL_SHAPES, _lib
    from yoho_b200 import dist as yd
    from yoho_b200.nn import init_encoder_params
    from yoho_b200.pipeline import ClipEncoder
    from yoho_b200.synth import synth_batch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: yoho_b200 has no CPU fallback (use --impl reference for the CPU oracle)")
    rank, local_rank, world = yd.init_process_group()
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # one process per GPU: keep each rank's threads and pinned staging buffers on the GPU's own NUMA node
    host_cpus = yd.bind_host_to_gpu(local_rank) if world > 1 else []
    lib = _lib.load()
    pk = peaks()

    model, clips, stages, desc = WORKLOADS[args.workload]
    if args.clips > 0:
        clips = args.clips
    n_mels, dims, heads, layers = MODEL_SHAPES[model]
    total_clips = clips * world
    lo, hi = yd.shard_range(total_clips, rank, world)      # this rank's contiguous clip range
    B = hi - lo

    params = init_encoder_params(n_mels, dims, layers if "encoder" in stages else 0, seed=42)
    enc = ClipEncoder((n_mels, dims, heads, layers if "encoder" in stages else 0), params, n_samples=N_SAMPLES,
                      norm=args.norm, compute="bf16", device=local_rank)
    pcm_host, out_host = enc.host_buffers(B)
    pcm_host.copy_(torch.from_numpy(synth_batch(B, N_SAMPLES, first_index=lo, unique=min(B, 16), mix=True)))
    pcm_dev = pcm_host.to(dev)
    mel_dev = torch.empty((B, enc.frames, n_mels), dtype=torch.float32, device=dev)
    out_dev = torch.empty((B, enc.seq_len, dims), dtype=torch.float32, device=dev) if "encoder" in stages else None
    if stages == "encoder":
        enc.logmel(pcm_dev, mel_dev)

    # one clip per call is launch-bound (~60 launches): replay the launch sequence from a CUDA graph (whisper.py:255-258 is this call)
    use_graph = stages == "logmel+encoder" and B <= 2 and not args.no_graph

    def step_device():
        if stages == "logmel":
            enc.logmel(pcm_dev, mel_dev)
        elif stages == "encoder":
            enc.encoder(mel_dev, out_dev)
        elif use_graph:
            enc.encode_graphed(pcm_dev, out_dev)
        else:
            enc.encode(pcm_dev, out_dev, chunk_clips=args.chunk or None)

    mel_host = torch.empty((B, enc.frames, n_mels), dtype=torch.float32, pin_memory=True) if stages != "logmel+encoder" else None

    def step_host():
        if stages == "logmel+encoder":
            enc.encode_host(pcm_host, out_host, chunk_clips=args.e2e_chunk)
        elif stages == "logmel":
            d = pcm_host.to(dev, non_blocking=True)
            enc.logmel(d, mel_dev)
            mel_host.copy_(mel_dev, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        else:
            mel_dev.copy_(mel_host, non_blocking=True)
            enc.encoder(mel_dev, out_dev)
            out_host.copy_(out_dev, non_blocking=True)
            torch.cuda.current_stream().synchronize()

    if stages == "encoder":
        mel_host.copy_(mel_dev)

    def timed(fn, steps, profile=False):
        yd.barrier()
        torch.cuda.synchronize()
        if profile:
            lib.yoho_prof_enable(1)
        n0 = lib.yoho_launch_count()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        yd.barrier()
        ms = a.elapsed_time(b)
        return yd.max_over_ranks(ms, dev), lib.yoho_launch_count() - n0

    # ---- device-resident metric ------------------------------------------------------------
    # `value`: exactly K steps with nothing but the kernels between the two events (clocks sampled beside it).  The
    # rooflines come from a second pass of K steps with the library's per-launch event records switched on: ~100 event
    # pairs per step keep consecutive kernels from overlapping their launch, which costs 2-4 % (`ms_per_step_profiled`).
    # warm-up: at least W steps AND at least ~2.5 s of back-to-back steps, so that the power-capped clock has settled before
    # the timed region (with 0.3 s of warm-up the first timed pass measured 2-3 % slower than an identical second pass)
    # The nvidia-smi sampler is started at the beginning of the warm-up (its start-up perturbs the kernels for about a second:
    # the first timed pass measured 2.7 % slower than an identical pass without it) and samples at 5 Hz through the timed
    # region; only the samples taken inside that region are reported.
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t_warm = time.perf_counter()
    n_warm = 0
    while n_warm < args.warmup or (time.perf_counter() - t_warm < 2.5 and n_warm < 4000):
        step_device()
        n_warm += 1
        if n_warm % 4 == 0:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    sampler.mark()
    ms_plain_total, launches = timed(step_device, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_plain_total / args.steps
    value = total_clips * CLIP_SECONDS / (ms_step * 1e-3)
    if use_graph:
        launches = getattr(enc, "graph_kernels", 0) * args.steps     # kernel nodes replayed from the captured graph
    ms_total, _ = timed(step_device, args.steps, profile=True)
    ncls = lib.yoho_prof_num_classes()
    p_ms, p_work, p_n = (C.c_double * ncls)(), (C.c_double * ncls)(), (C.c_longlong * ncls)()
    _lib.check(lib.yoho_prof_collect(p_ms, p_work, p_n, ncls))
    lib.yoho_prof_enable(0)
    ms_step_profiled = ms_total / args.steps

    # ---- end-to-end metric (host buffers, copies inside the timed region) ---------------------
    for _ in range(2):
        step_host()
    ms_e2e_total, _ = timed(step_host, args.steps)
    ms_e2e = ms_e2e_total / args.steps
    e2e_value = total_clips * CLIP_SECONDS / (ms_e2e * 1e-3)
    h2d = B * (N_SAMPLES * 4 if stages != "encoder" else enc.frames * n_mels * 4)
    d2h = B * (enc.seq_len * dims * 4 if "encoder" in stages else enc.frames * n_mels * 4)

    # ---- second host-facing line: int16 PCM in (what ffmpeg delivers, audio.py:11-18), bf16 states out: half the bytes both ways
    e2e_compact = None
    if stages == "logmel+encoder" and B >= 2:
        enc_c = ClipEncoder((n_mels, dims, heads, layers), params, n_samples=N_SAMPLES, norm=args.norm, compute="bf16",
                            device=local_rank, states_dtype="bf16")
        pcm16_host, out16_host = enc_c.host_buffers(B, torch.int16)
        pcm16_host.copy_((pcm_host * 32767.0).round().to(torch.int16))
        for _ in range(2):
            enc_c.encode_host(pcm16_host, out16_host, chunk_clips=args.e2e_chunk)
        ms_c_total, _ = timed(lambda: enc_c.encode_host(pcm16_host, out16_host, chunk_clips=args.e2e_chunk), args.steps)
        ms_c = ms_c_total / args.steps
        e2e_compact = {"value": round(total_clips * CLIP_SECONDS / (ms_c * 1e-3), 1), "unit": "x realtime", "ms_per_step": round(ms_c, 4),
                       "h2d_bytes_per_step": B * N_SAMPLES * 2, "d2h_bytes_per_step": B * enc.seq_len * dims * 2,
                       "note": "int16 PCM in, bf16 encoder states out (yoho_logmel_s16, out_dtype = YOHO_DTYPE_BF16); the headline e2e stays float32 both ways"}
        del enc_c, pcm16_host, out16_host

    # ---- optional NCCL all-gather of the encoder states, off the hot path (reported as a separate line) ----
    allgather = None
    if world > 1 and out_dev is not None:
        try:
            import torch.distributed as tdist
            allgather = {}
            for name, src in (("f32", out_dev), ("bf16", out_dev.to(torch.bfloat16))):
                dst = torch.empty((world * B,) + tuple(src.shape[1:]), dtype=src.dtype, device=dev)
                for _ in range(2):
                    tdist.all_gather_into_tensor(dst, src)
                reps = 5
                ms_g, _ = timed(lambda: tdist.all_gather_into_tensor(dst, src), reps)
                ms_g /= reps
                nbytes = dst.numel() * dst.element_size()
                allgather[name] = {"ms": round(ms_g, 3), "bytes_gathered_per_rank": nbytes,
                                   "algbw_gbs": round(nbytes / ms_g / 1e6, 1), "busbw_gbs": round(nbytes * (world - 1) / world / ms_g / 1e6, 1),
                                   "share_of_step": round(ms_g / ms_step, 4)}
                del dst
            allgather["note"] = f"torch.distributed.all_gather_into_tensor (NCCL) of ({B}, {enc.seq_len}, {dims}) states per rank over {world} ranks; not inside `value` / `e2e`"
        except Exception as exc:
            allgather = {"error": repr(exc)}

    # ---- rooflines from the per-class CUDA events of the timed region -----------------------------
    classes = {lib.yoho_prof_class_name(i).decode(): (p_ms[i], p_work[i], p_n[i]) for i in range(ncls)}
    rooflines = {}
    for name, (cms, work, n) in classes.items():
        if n == 0 or cms <= 0:
            continue
        tensor = name in ("gemm_tcgen05", "attention_tcgen05")
        achieved = work / (cms * 1e-3) / (1e12 if tensor else 1e9)
        peak = pk["bf16_sustained"] if tensor else pk["hbm_gbs"]
        rooflines[name] = {"bound": "tensor" if tensor else "hbm", "achieved": round(achieved, 2), "peak": peak,
                           "unit": "TFLOP/s" if tensor else "GB/s", "frac": round(achieved / peak, 4),
                           "launches_per_step": n / args.steps, "ms_per_step": round(cms / args.steps, 4),
                           "share_of_step": round(cms / ms_total, 4)}
    dominant = max(rooflines, key=lambda k: rooflines[k]["ms_per_step"]) if rooflines else None
    # DRAM bytes per launch of the dominant kernel, from the committed `ncu --set full` capture (profiles/traffic.json,
    # written by tools/ncu_traffic.py from a 32-clip run of the same kernels), scaled to this run's clips per launch.
    traffic, traffic_note = None, "no profiles/traffic.json"
    try:
        tj = json.loads((ROOT / "profiles" / "traffic.json").read_text())
        if dominant in tj:
            per_launch_clips = B if dominant != "logmel" else B / max(1.0, rooflines[dominant]["launches_per_step"])
            traffic = round(tj[dominant]["bytes_per_launch_per_clip"] * per_launch_clips)
            traffic_note = (f"ncu dram__bytes_read.sum + dram__bytes_write.sum, mean of {tj[dominant]['launches']} launch(es) at "
                            f"{tj[dominant]['clips']} clips, scaled to {per_launch_clips:.0f} clips per launch")
    except Exception as exc:
        traffic_note = f"profiles/traffic.json unreadable: {exc!r}"
    for name, r in rooflines.items():
        if r["bound"] == "hbm" and r["frac"] > 1.0:
            r["note"] = ("algorithmic bytes over the measured copy bandwidth: above 1 because part of the kernel's input was written by the kernel "
                         "before it and is still in the 126 MB L2 (measured DRAM traffic is below the algorithmic bytes)")
    if dominant == "gemm_tcgen05" and "encoder" in stages and os.environ.get("YOHO_LN_FUSE", "1")[:1] != "0":
        rooflines[dominant]["note"] = ("FLOP of the GEMMs over a time that also holds the work of the 2 LayerNorms per block folded into their epilogues "
                                       "(residual rows fetched and re-stored, bf16 copy, row statistics); A/B on one box, profiles/r02_lnfuse.txt: "
                                       "YOHO_LN_FUSE=0 gives gemm 1295 TFLOP/s (0.90) + 7.4 ms of LayerNorm kernels and a step 1.8 % slower")
    roofline = dict(rooflines[dominant], kernel=dominant, traffic=traffic, traffic_note=traffic_note,
                    peak_source=f"{pk['source']} ({'sustained bf16 cuBLAS' if rooflines[dominant]['bound'] == 'tensor' else 'copy bandwidth'}, MEASURED_PEAKS.json)") if dominant else None

    if rank != 0:
        return 0
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            v, cores, sample, _ = cpu_oracle_rtfx(model, stages)
            cpu = {"value": round(v, 2), "unit": "x realtime", "cores": cores, "kind": "port", "sample": sample}
        except Exception as exc:  # the GPU numbers stand on their own
            cpu = {"value": None, "unit": "x realtime", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {exc!r}"}
    line = {
        "metric": "audio-sec/sec (RTFx) log-mel+encoder" if stages == "logmel+encoder" else f"audio-sec/sec (RTFx) {stages}",
        "value": round(value, 1), "unit": "x realtime", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_step, 4), "ms_per_step_profiled": round(ms_step_profiled, 4), "warmup_steps_run": n_warm, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if "encoder" in stages else "f32", "data": "synthetic",
        "config": {"workload": desc, "model": model, "clips_per_gpu": clips, "global_clips": total_clips, "clip_seconds": CLIP_SECONDS,
                   "stages": stages, "norm": args.norm, "compute": "bf16 operands / fp32 accumulate (tcgen05); fp32 frontend, residual stream and LayerNorm",
                   "parallelism": f"clips sharded over {world} GPU(s), no collective", "host_cpus_per_rank": len(host_cpus) or None,
                   "weights": "random-init (seed 42)", "cuda_graph": bool(use_graph),
                   "layernorm": ("folded into the GEMM epilogues on either side (gemm_tcgen05 time includes that work; only the final LayerNorm is a kernel)"
                                 if os.environ.get("YOHO_LN_FUSE", "1")[:1] != "0" else "stand-alone kernels (YOHO_LN_FUSE=0)"),
                   "clips": f"{B} distinct synthetic clips per rank: 16 synthesised from scratch, the others each a mixture of two of them with a clip-specific rotation (yoho_b200.synth.synth_batch(mix=True))",
                   "l2": f"inputs larger than L2: {B * N_SAMPLES * 4 / 1e6:.0f} MB PCM per GPU per step, activations stream through HBM"},
        "e2e": {"value": round(e2e_value, 1), "unit": "x realtime", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(ms_e2e, 4), "chunk_clips": getattr(enc, "last_chunk_clips", args.e2e_chunk)},
        "e2e_compact": e2e_compact, "allgather": allgather,
        "gpu_launches": int(launches),
        "roofline": roofline, "rooflines": rooflines,
        "tflops_encoder": round(total_clips * GFLOP_PER_CLIP[model] / ms_step / world, 2) if "encoder" in stages else None,
        "clocks": clocks, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    rc = main()
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass
    sys.exit(rc)
