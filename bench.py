#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200 Q4_0 hot path.

    python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

Workload at N=1 (BASELINE.json configs[1]): Whisper Medium encoder, one 30 s clip per step, mel
[1,80,3000], random-init Q4_0 weights.  A "step" is one encoder pass over one batch of synthetic log-mel
input (batch = --batch clips per GPU, default 1).  metric = encoder 30 s-clips/s, whole job.

  value     clips/s with the mel already resident in HBM (CUDA-graph replay of the encoder)
  e2e       clips/s through the public API from HOST buffers: pinned mel -> H2D -> encode -> D2H of the
            encoder output, every step
  roofline  the dominant kernel (tcgen05 Q4 GEMM): algorithmic flop / CUDA-event time of its launches
            inside the step, against the measured bf16 peak of MEASURED_PEAKS.json
  cpu_baseline  the reference's CPU dequant+matmul (C port, oracle/) on the host cores, bounded sample

Multi-GPU: utterances are independent, so each rank encodes its own clips (no collective on the data
path); value = clips of all ranks / max-over-ranks time; scaling "weak".

The same JSON line carries a `legs` object measured in the same process after the headline, one entry per remaining
BASELINE.json config / north_star target (all on Whisper Large-V3 shapes unless stated):
  gemv_1280x1280        config 1: Q4_0 dequant+matmul 1280x1280 at M = 1, GPU GEMV (HBM GB/s) beside the CPU port
  decode_step           config 3: one greedy decode step at batch 1, kv_len in {4, 224, 447}, CUDA-graph replay:
                        tokens/s and HBM GB/s over the bytes the step reads (Q4 weights + f16 K/V + f16 logits matrix)
  encoder_large_v3_b8   the GEMM roofline at the shape north_star quotes it on (12 000 rows), fraction of BURST peak
  transcribe_256        config 5: 256 synthetic clips through the full transcribe path, split contiguously over the
                        ranks (STRONG scaling: the clip count is fixed), clips/s and tokens/s; runs at every N
(--legs none skips them; at N > 1 only transcribe_256 runs.)
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "encoder 30s-clips/s"
UNIT = "clips/s"


# ---------------------------------------------------------------------------------------- helpers
def shard_clips(total: int, rank: int, world: int):
    """Contiguous split of `total` utterances over `world` ranks."""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def max_over_ranks(x: float) -> float:
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(x)
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"tflops": float(d.get("bf16_tflops", 1590.0)), "tflops_sustained": float(d.get("bf16_tflops_sustained", 1400.0)),
                "hbm_gbs": float(d.get("hbm_gbs", 6650.0)), "source": "measured"}
    return {"tflops": 1590.0, "tflops_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback"}


class ClockSampler:
    """SM clock and throttle reasons sampled through NVML every 5 ms DURING the timed region
    (B200_PROFILING.md clocks line; NVML is what nvidia-smi reads, without its process start-up delay)."""
    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index: int):
        self.index, self.clocks, self.power, self.reasons, self.mx = index, [], [], set(), None
        self.stop_flag, self.thr, self.err = threading.Event(), None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[self.index]) if visible and all(v.strip().isdigit() for v in visible.split(",")) else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.thr = threading.Thread(target=self._poll, daemon=True)
            self.thr.start()
        except Exception as e:          # noqa: BLE001 - sampling is best effort, the bench itself must still run
            self.err = f"{type(e).__name__}: {e}"

    def _poll(self):
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                self.clocks.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for name, bit in self.BAD.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception as e:      # noqa: BLE001
                self.err = f"{type(e).__name__}: {e}"
                return
            time.sleep(0.005)

    def stop(self):
        self.stop_flag.set()
        if self.thr:
            self.thr.join(timeout=2)
        if not self.clocks:
            return {"sm_mhz": None, "sm_max_mhz": self.mx, "samples": 0, "reasons": [f"nvml unavailable ({self.err})"]}
        return {"sm_mhz": statistics.median(self.clocks), "sm_min_mhz": min(self.clocks), "sm_max_mhz": self.mx,
                "power_w": round(statistics.median(self.power), 1), "samples": len(self.clocks), "reasons": sorted(self.reasons)}


def ncu_traffic(model: str, batch: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of the Q4 GEMM kernel, mean per launch over one step, from the
    committed ncu pass of this command (profiles/traffic.json, written by scripts/ncu_traffic.py)."""
    p = ROOT / "profiles" / "traffic.json"
    if not p.exists():
        return None, "no ncu capture committed"
    try:
        d = json.loads(p.read_text()).get(f"{model}_b{batch}")
        return (float(d["gemm_bytes_per_launch"]), d.get("source", "profiles/traffic.json")) if d else (None, "no capture for this workload")
    except (ValueError, KeyError, TypeError):
        return None, "profiles/traffic.json unreadable"


def preset(name: str):
    import whisper_burn_b200 as wb
    return wb.WhisperConfig.medium() if name == "medium" else wb.WhisperConfig.large_v3()


def encoder_q4_flops(cfg, clips: int) -> float:
    """2*M*N*K over the six Q4 GEMMs of every encoder layer, M = 1500 per clip."""
    D, F, M = cfg.n_audio_state, 4 * cfg.n_audio_state, cfg.n_audio_ctx
    return clips * cfg.n_audio_layer * 2.0 * M * (4 * D * D + 2 * D * F)


# ---------------------------------------------------------------------------------- reference arm
def cpu_sample(cfg, layers: int, threads: int):
    """The reference's CPU path for this workload: fused dequant+matmul (CPU twin of shader.wgsl:72-89,
    oracle/q4_oracle.c) for the six Q4 GEMMs of `layers` encoder layers at M = 1500, all host threads.
    Returns seconds for the sample."""
    import numpy as np
    from oracle import q4
    q4.set_threads(threads)
    rng = np.random.default_rng(0)
    D, F, M = cfg.n_audio_state, 4 * cfg.n_audio_state, cfg.n_audio_ctx
    shapes = [(D, D)] * 4 + [(F, D), (D, F)]
    ws = [(n, k, q4.quantize_q4_0_converter((rng.standard_normal(n * k) * 0.02).astype(np.float32))) for n, k in shapes]
    xs = {k: rng.standard_normal((1, M, k)).astype(np.float32) for k in (D, F)}
    t0 = time.perf_counter()
    for _ in range(layers):
        for n, k, qb in ws:
            q4.q4_matmul(xs[k], qb, n)
    return time.perf_counter() - t0


def run_reference(args):
    """The reference's CPU path for the workload's Q4 linears (C port of shader.wgsl:72-89 / tensor.rs:96-109, oracle/q4_oracle.c;
    the Rust crate cannot be built here) on all host threads.  Each step is a bounded sample -- the six Q4 GEMMs of `layers`
    encoder layers at M = 1500, sized to ~3 s -- and is reported as such: `ms_per_step` is the measured time of one sample step,
    `steps` the number actually timed, so steps x ms_per_step is the wall time of the timed region; `value` scales the sample to
    the whole clip (conv / attention / LayerNorm not charged to the CPU)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import q4
    cfg = preset(args.model)
    threads = q4.host_cores()
    t1 = cpu_sample(cfg, 1, threads)                       # sizing run: one layer
    layers = max(1, min(cfg.n_audio_layer, int(round(3.0 / max(t1, 1e-3)))))
    frac = layers / cfg.n_audio_layer
    for _ in range(max(0, min(args.warmup, 2))):
        cpu_sample(cfg, layers, threads)
    steps = max(1, min(args.steps, int(150.0 / max(t1 * layers, 1e-3))))     # the whole run ends within a few minutes
    times = [cpu_sample(cfg, layers, threads) for _ in range(steps)]
    t = sum(times) / len(times)
    clips_per_s = frac / t
    sample = (f"fused Q4_0 dequant+matmul of the 6 linears of {layers} of {cfg.n_audio_layer} encoder layers at M=1500 "
              f"({frac * encoder_q4_flops(cfg, 1) / 1e9:.1f} GFLOP) per step, scaled x{cfg.n_audio_layer / layers:.2f} to the clip; "
              f"conv/attention/LN not charged")
    line = {"impl": "reference", "metric": METRIC, "value": clips_per_s, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": max(0, min(args.warmup, 2)), "ms_per_step": t * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"whisper-{args.model} encoder, 1x30s mel [1,{cfg.n_mels},{2 * cfg.n_audio_ctx}] per GPU per step, "
                                   f"random-init Q4_0 weights", "batch_per_gpu": 1,
                       "arm": "reference's CPU path for the Q4 linears (C port, oracle/q4_oracle.c) on all host threads; each step is a "
                              f"bounded sample of {layers} of {cfg.n_audio_layer} encoder layers; ms_per_step is the sample's own time, "
                              "value is scaled to the whole clip",
                       "sample_fraction_of_clip": frac},
            "cpu_baseline": {"value": clips_per_s, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": clips_per_s, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0



# --------------------------------------------------------------------------------------------- legs
def _instrumented_encoder(L, model, mel_dev, step_s, prof_steps=5):
    """Per-kernel CUDA-event timing of the encoder step: every instrumented step is queued behind a GPU-side spin that
    outlasts the host's enqueue of the whole step, so an event pair brackets the kernel alone."""
    import ctypes as C
    import torch
    import whisper_burn_b200 as wb
    kinds = {"q4_gemm": 0, "conv_gemm": 1, "gemv": 2, "attention": 3, "layernorm": 4}
    L.wq4_profile_read.argtypes = [C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.wq4_profile_reset()
    L.wq4_profile_enable(1)
    spin_cycles = int(max(8e-3, 1.5 * step_s) * 1.9e9)
    for _ in range(prof_steps):
        torch.cuda._sleep(spin_cycles)
        model.encode(mel_dev, want_output=False)
        torch.cuda.synchronize()
    L.wq4_profile_enable(0)
    prof = {}
    for name, k in kinds.items():
        n, ms, work = C.c_int64(), C.c_double(), C.c_double()
        wb._lib.check(L.wq4_profile_read(k, C.byref(n), C.byref(ms), C.byref(work)))
        prof[name] = {"launches_per_step": n.value // prof_steps, "ms_per_step": ms.value / prof_steps,
                      "work_per_step": work.value / prof_steps}
    L.wq4_profile_reset()
    return prof


def _in_step(prof, step_ms, g, peak):
    tot = sum(v["ms_per_step"] for v in prof.values())
    if not tot or not g["ms_per_step"]:
        return None
    ms = g["ms_per_step"] * step_ms / tot
    ach = g["work_per_step"] / (ms * 1e-3) / 1e12
    return {"kernel_ms_per_step": ms, "achieved": ach, "frac": ach / peak, "sum_of_isolated_kernel_ms": tot, "step_ms": step_ms}


def _graph_time(stream, fn, reps, warm=3):
    """Seconds per call of fn() under CUDA-graph replay on `stream` (CUDA events, synchronised on both sides)."""
    import torch
    for _ in range(warm):
        fn()
    stream.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        fn()
    for _ in range(warm):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        g.replay()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def leg_gemv(dev, stream, peaks, no_cpu):
    """BASELINE.json config 1: Q4_0 dequant + matmul 1280x1280 (tests.rs:412-480 generators) at M = 1: the CUDA-core GEMV
    rotating through 160 distinct copies of the weight (147 MB > L2, so every launch streams from HBM), beside the CPU
    reference (dequantise then matmul, tensor.rs:96-109 + tests.rs:172-184, and its fused twin) on the host cores."""
    import ctypes as C
    import numpy as np
    import torch
    import whisper_burn_b200 as wb
    from oracle import q4
    n = k = 1280
    w = (0.05 * np.cos(0.0007 * np.arange(n * k, dtype=np.float32))).astype(np.float32)
    x = (0.1 * np.sin(0.001 * np.arange(k, dtype=np.float32))).astype(np.float32).reshape(1, 1, k)
    qb = q4.quantize_q4_0_tests(w)
    copies = [wb.Q4Tensor.from_q4_bytes(qb, [n, k], dev) for _ in range(160)]
    xd = torch.from_numpy(x).to(dev)
    yd = torch.empty((1, 1, n), dtype=torch.float32, device=dev)
    L = wb.lib()
    sp = C.c_void_p(stream.cuda_stream)

    def run():
        for t in copies:
            wb._lib.check(L.wq4_linear_forward_ex(t._h, None, C.c_void_p(xd.data_ptr()), None, C.c_void_p(yd.data_ptr()), None, None,
                                                  1, 0, None, 0, sp))
    dt = _graph_time(stream, run, reps=20) / len(copies)
    bytes_ = n * k // 32 * 18 + 4 * k + 4 * n
    leg = {"config": "Q4_0 dequant+matmul 1280x1280, M=1 (BASELINE.json configs[0])", "us_per_launch": dt * 1e6,
           "algorithmic_bytes": bytes_, "roofline": {"bound": "hbm", "achieved": bytes_ / dt / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                                     "frac": bytes_ / dt / 1e9 / peaks["hbm_gbs"]},
           "note": "a single 0.9 MB GEMV is launch-latency bound (0.15 us of streaming); the decode_step leg is the number that matters"}
    if not no_cpu:
        threads = q4.host_cores()
        q4.set_threads(threads)
        reps, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < 0.5:
            q4.q4_matmul(x, qb, n)
            reps += 1
        fused = (time.perf_counter() - t0) / reps
        reps, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < 0.5:
            q4.dequant_then_matmul(x, qb, 1, k, n)
            reps += 1
        two = (time.perf_counter() - t0) / reps
        leg["cpu_baseline"] = {"fused_us": fused * 1e6, "dequant_then_matmul_us": two * 1e6, "cores": threads, "kind": "port",
                               "sample": "oracle/q4_oracle.c on all host threads, 0.5 s of repeats each"}
    return leg


def leg_decode(model, cfg, dev, stream, peaks):
    """BASELINE.json config 3: Whisper Large-V3 greedy decoder step, batch 1, KV cache in place, kv_len in {4, 224, 447}."""
    import numpy as np
    import torch
    import whisper_burn_b200 as wb
    D, F, Ld, V, T = model.d_model, model.ffn, cfg.n_text_layer, model.vocab, cfg.n_audio_ctx
    q4_bytes = Ld * (6 * D * D + 2 * D * F) // 32 * 18          # self q,k,v,out + cross q,out + fc1 + fc2: 412.9 MB at Large-V3
    rng = np.random.default_rng(0)
    mel = torch.from_numpy(np.clip(rng.standard_normal((1, cfg.n_mels, 2 * T)).astype(np.float32) * 0.5, -1, 1.5)).to(dev)
    model.encode(mel, want_output=False)
    model.decoder_pass([[50258, 50259, cfg.transcribe_token(), cfg.no_timestamps_token()]], 0, 0, recompute_cross=True)
    out = {}
    pos = 4
    for kv in (4, 224, 447):
        while pos < kv:                                  # fill the self-attention cache with real rows up to kv_len
            model.decode_step_device(1, pos, pos)
            pos += 1
        wb.reset_launch_count()
        model.decode_step_device(1, kv, kv)
        launches = wb.launch_count()
        dt = _graph_time(stream, lambda: model.decode_step_device(1, kv, kv), reps=100)
        kv_bytes = Ld * (T + kv + 1) * D * 2 * 2       # f16 K and V rows read: cross (1500) + self (kv_len + the new row)
        logit_bytes = V * D * 2                          # the f16 logits matrix (the f32 embedding is only read for one row)
        total = q4_bytes + kv_bytes + logit_bytes
        out[f"kv{kv}"] = {"us_per_step": dt * 1e6, "tokens_per_s": 1.0 / dt, "kernel_launches_per_step": launches,
                          "bytes": {"q4_weights": q4_bytes, "kv_f16": kv_bytes, "logits_f16": logit_bytes},
                          "roofline": {"bound": "hbm", "achieved": total / dt / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                       "frac": total / dt / 1e9 / peaks["hbm_gbs"], "q4_only_GBps": q4_bytes / dt / 1e9,
                                       "q4_only_frac": q4_bytes / dt / 1e9 / peaks["hbm_gbs"]}}
    out["config"] = "whisper-large_v3 greedy decode step, batch 1, 32 layers, CUDA-graph replay (BASELINE.json configs[2])"
    out["roofline"] = out["kv224"]["roofline"]
    return out


def leg_encoder_b8(L, model, cfg, dev, stream, peaks):
    """The encoder GEMM roofline at the shape north_star quotes it on: Large-V3, 8 clips per step (12 000 rows)."""
    import numpy as np
    import torch
    rng = np.random.default_rng(5)
    B = 8
    mel = torch.from_numpy(np.clip(rng.standard_normal((B, cfg.n_mels, 2 * cfg.n_audio_ctx)).astype(np.float32) * 0.5, -1, 1.5)).to(dev)
    dt = _graph_time(stream, lambda: model.encode(mel, want_output=False), reps=10)
    prof = _instrumented_encoder(L, model, mel, dt, prof_steps=3)
    g = prof["q4_gemm"]
    achieved = g["work_per_step"] / (g["ms_per_step"] * 1e-3) / 1e12 if g["ms_per_step"] > 0 else 0.0
    return {"config": "whisper-large_v3 encoder, 8x30s clips per step (M = 12000)", "ms_per_step": dt * 1e3, "clips_per_s": B / dt,
            "roofline": {"bound": "tensor", "kernel": "q4_chain_kernel + q4_gemm2_kernel", "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["tflops"], "frac_of_sustained_peak": achieved / peaks["tflops_sustained"],
                         "launches_per_step": g["launches_per_step"], "kernel_ms_per_step": g["ms_per_step"],
                         "algorithmic_flop_per_step": g["work_per_step"]},
            "breakdown_ms_per_step": {k: round(v["ms_per_step"], 4) for k, v in prof.items()}}


def leg_transcribe(model, cfg, dev, rank, world, clips=256, micro=256):
    """BASELINE.json config 5: 256 synthetic 30 s clips (seeds 100..355) through the full transcribe path, split contiguously
    over the ranks; no collective on the data path.  Strong scaling: the clip count is fixed.  A rank decodes all its clips
    as ONE micro-batch (256 clips on one GPU: 63 GB of f16 cross K/V + 19 GB of self K/V + ~17 GB of encoder activations out
    of 180 GB): a decode step reads the Q4 weights once whatever the batch, so 256 clips per step give 16 100 tokens/s
    against 10 600 at 64 (71.9 against 47.3 clips/s, measured)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    lo, hi = shard_clips(clips, rank, world)
    mine = hi - lo
    micro = max(1, min(micro, mine))
    model.reserve(micro)
    host = torch.empty((mine, cfg.n_mels, 2 * cfg.n_audio_ctx), dtype=torch.float32).pin_memory()
    for i in range(mine):
        rng = np.random.default_rng(100 + lo + i)
        host[i] = torch.from_numpy(np.clip(rng.standard_normal((cfg.n_mels, 2 * cfg.n_audio_ctx)).astype(np.float32) * 0.5, -1.0, 1.5))
    model.transcribe(host[:min(2, mine)].to(dev), "en", max_new_tokens=8)          # warm-up: module load, descriptors, plans
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    n_tok = 0
    for s in range(0, mine, micro):
        mel = host[s:s + micro].to(dev, non_blocking=True)          # H2D inside the timed region
        ids, _ = model.transcribe(mel, "en")                        # token ids come back to the host
        n_tok += sum(len(t) for t in ids)
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    if world > 1:
        dist.barrier()
    dt = max_over_ranks(e0.elapsed_time(e1) * 1e-3)
    wall = max_over_ranks(wall)
    tok = torch.tensor([n_tok], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tok)
    return {"config": f"whisper-large_v3 full transcribe of {clips} synthetic 30s clips, utterance split x{world}, micro-batch {micro} "
                      f"(BASELINE.json configs[4])", "scaling": "strong", "clips": clips, "clips_per_gpu": mine, "micro_batch": micro,
            "seconds": dt, "wall_seconds": wall, "clips_per_s": clips / dt, "tokens_per_s": float(tok.item()) / dt,
            "tokens": int(tok.item()), "timing": "CUDA events around the whole loop (H2D of the mels and D2H of the ids inside), max over ranks"}


def run_legs(args, L, dev, stream, rank, world, local, peaks):
    import gc
    import torch
    import whisper_burn_b200 as wb
    from whisper_burn_b200.gguf_file import random_whisper_gguf
    legs = {}
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        if world == 1:
            legs["gemv_1280x1280"] = leg_gemv(dev, stream, peaks, args.no_cpu)
        cfg = wb.WhisperConfig.large_v3()
        model = wb.load_whisper_from_gguf(random_whisper_gguf(cfg, seed=0), cfg, dev)
        if world == 1:
            model.reserve(1)
            legs["decode_step"] = leg_decode(model, cfg, dev, stream, peaks)
            legs["encoder_large_v3_b8"] = leg_encoder_b8(L, model, cfg, dev, stream, peaks)
        legs["transcribe_256"] = leg_transcribe(model, cfg, dev, rank, world)
        del model
        gc.collect()
    legs["seconds_spent"] = round(time.perf_counter() - t0, 1)
    return legs


# ------------------------------------------------------------------------------------------ own arm
def run_own(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import ctypes as C
    import whisper_burn_b200 as wb
    from whisper_burn_b200.gguf_file import random_whisper_gguf

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = wb.lib()
    cfg = preset(args.model)
    B = args.batch
    # encoder-only checkpoint: the decoder stack is not part of this workload, so it is neither generated nor loaded
    import dataclasses
    blob = random_whisper_gguf(cfg, seed=0, with_decoder=False)
    model = wb.load_whisper_from_gguf(blob, dataclasses.replace(cfg, n_text_layer=0), dev)
    del blob
    model.reserve(B)
    rng = np.random.default_rng(100 + rank)
    mel_host = torch.from_numpy(np.clip(rng.standard_normal((B, cfg.n_mels, 2 * cfg.n_audio_ctx)).astype(np.float32) * 0.5, -1, 1.5)).pin_memory()
    mel_dev = mel_host.to(dev)
    out_host = torch.empty((B, cfg.n_audio_ctx, cfg.n_audio_state), dtype=torch.float32).pin_memory()

    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        for _ in range(max(3, args.warmup)):
            model.encode(mel_dev, want_output=False)
        stream.synchronize()
        wb.reset_launch_count()
        model.encode(mel_dev, want_output=False)
        launches_per_step = wb.launch_count()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            model.encode(mel_dev, want_output=False)
        for _ in range(max(3, args.warmup)):
            graph.replay()
        stream.synchronize()

        # ---- value: device-resident inputs, K steps, CUDA events, barrier + sync on both sides
        sampler = ClockSampler(local)
        sampler.start()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            graph.replay()
        e1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        dt = max_over_ranks(e0.elapsed_time(e1) * 1e-3)
        clocks = sampler.stop()

        # ---- e2e: host buffers through the public API, copies inside the timed region
        for _ in range(3):
            model.encode(mel_host.to(dev, non_blocking=True))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        # Every step copies its own mel from pinned host memory and reads its own result back; the read-back runs on
        # a second stream (double-buffered pinned destination) so it overlaps the next step's encode, as a serving
        # loop would do.  The host only blocks on the read-back of step i-2 (buffer reuse) and at the end.
        copy_stream = torch.cuda.Stream(device=dev)
        out_bufs = [out_host, torch.empty_like(out_host).pin_memory()]

        def e2e_pass(n_steps):
            done = [None, None]
            for i in range(n_steps):
                m = mel_host.to(dev, non_blocking=True)
                y = model.encode(m)
                ready = torch.cuda.Event()
                ready.record(stream)
                if done[i % 2] is not None:
                    done[i % 2].synchronize()
                copy_stream.wait_event(ready)
                with torch.cuda.stream(copy_stream):
                    out_bufs[i % 2].copy_(y, non_blocking=True)
                    y.record_stream(copy_stream)
                    done[i % 2] = torch.cuda.Event()
                    done[i % 2].record(copy_stream)
            torch.cuda.synchronize()

        e2e_pass(min(args.steps, 5))          # untimed: first touch of the second pinned buffer and of the copy stream
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        e2e_pass(args.steps)
        dt_e2e = max_over_ranks(time.perf_counter() - t0)

        # ---- roofline: per-kernel CUDA-event timing of the same step (instrumented pass)
        prof = _instrumented_encoder(L, model, mel_dev, dt / args.steps)

    peaks = measured_peaks()
    # The GEMM launches are timed by event pairs in an instrumented pass of isolated steps (each behind a GPU-side spin), i.e.
    # a kernel timed alone: the denominator is the BURST measured bf16 rate; the sustained figure is kept beside it.
    peak = peaks["tflops"]
    traffic, traffic_src = ncu_traffic(args.model, B)
    g = prof["q4_gemm"]
    achieved = g["work_per_step"] / (g["ms_per_step"] * 1e-3) / 1e12 if g["ms_per_step"] > 0 else 0.0
    step_ms = dt / args.steps * 1e3
    errflag = L.wq4_device_error_flag(local)
    line = {
        "metric": METRIC, "value": world * B * args.steps / dt, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": f"whisper-{args.model} encoder, {B}x30s mel [{B},{cfg.n_mels},{2 * cfg.n_audio_ctx}] per GPU per step, "
                               f"random-init Q4_0 weights", "batch_per_gpu": B, "parallelism": f"utterance split x{world}, no collective",
                   "l2": f"inputs larger than L2: {model.weight_bytes / 1e6:.0f} MB of weights streamed per step vs 126 MB L2",
                   "operands": "f16 tensor-core operands, f32 accumulate (bf16 misses the 1e-3 bar, DESIGN.md)",
                   "timing": "CUDA events around K CUDA-graph replays, max over ranks"},
        "clocks": clocks,
        "e2e": {"value": world * B * args.steps / dt_e2e, "unit": UNIT,
                "h2d_bytes_per_step": int(mel_host.numel() * 4), "d2h_bytes_per_step": int(out_host.numel() * 4)},
        "gpu_launches": int(launches_per_step * args.steps),
        "roofline": {"bound": "tensor", "kernel": "q4_chain_kernel + q4_gemm2_kernel (tcgen05 cta_group::2, Q4_0 dequant -> TMEM, f16 MMA; one chained launch per encoder block: out-proj, fc1, fc2, next q|k|v with the LayerNorms folded in)",
                     "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                     "traffic_source": traffic_src,
                     "peak_source": f"{peaks['source']} burst bf16 (MEASURED_PEAKS.json); sustained {peaks['tflops_sustained']}",
                     "frac_of_sustained_peak": achieved / peaks["tflops_sustained"],
                     "launches_per_step": g["launches_per_step"], "kernel_ms_per_step": g["ms_per_step"],
                     "algorithmic_flop_per_step": g["work_per_step"],
                     "share_of_step": g["ms_per_step"] / sum(v["ms_per_step"] for v in prof.values()) if g["ms_per_step"] else None,
                     # The event pairs serialise launches that overlap under programmatic dependent launch, so the per-kernel
                     # times add up to MORE than the step they were taken from.  in_step attributes the measured step time to
                     # the kernels in proportion to those times (parts sum to the whole); `achieved` / `frac` above stay the
                     # conservative per-launch figures.
                     "in_step": _in_step(prof, step_ms, g, peak)},
        "breakdown_ms_per_step": {k: round(v["ms_per_step"], 4) for k, v in prof.items()},
        "device_error_flag": errflag,
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import q4
        threads = q4.host_cores()
        t1 = cpu_sample(cfg, 1, threads)                   # warm-up and sizing
        n_l = max(1, min(cfg.n_audio_layer, int(round(12.0 / max(t1, 1e-3)))))   # ~12 s of CPU work
        t = cpu_sample(cfg, n_l, threads)
        frac = n_l / cfg.n_audio_layer
        line["cpu_baseline"] = {"value": frac / t, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"fused Q4_0 dequant+matmul (oracle/q4_oracle.c, shader.wgsl:72-89 twin) of the 6 linears of {n_l} of "
                                          f"{cfg.n_audio_layer} encoder layers at M=1500, {t:.2f} s on {threads} threads, scaled x"
                                          f"{cfg.n_audio_layer / n_l:.2f}; conv/attention/LN not charged to the CPU"}
    if args.legs != "none":
        del model, graph
        line["legs"] = run_legs(args, L, dev, stream, rank, world, local, peaks)
        errflag = L.wq4_device_error_flag(local)
        line["device_error_flag"] = errflag
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0 if errflag == 0 else 3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--model", default="medium", choices=["medium", "large_v3"])
    ap.add_argument("--batch", type=int, default=1, help="clips per GPU per step")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--legs", default="all", choices=["all", "none"], help="extra legs: decode step, GEMV, Large-V3 x8 encoder, 256-clip transcribe")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_own(args)


if __name__ == "__main__":
    sys.exit(main())
