#!/usr/bin/env python3
"""Config 5 (BASELINE.json): Whisper Large-V3, a batch of 256 synthetic 30 s clips sharded by utterance over the
GPUs of one box; every clip goes through the full transcribe path (encoder -> cross K/V -> greedy decode).

    python scripts/bench_transcribe.py                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_transcribe.py --gpus N

Each rank loads the same random-init Q4_0 checkpoint (seed 0), takes the contiguous share of the clips that
bench.shard_clips gives it (clip i is generated from seed 100+i) and transcribes it in micro-batches of
--micro clips.  No collective runs on the data path; torch.distributed only carries the barrier and the max-over-ranks
of the device time.  With random-init weights EOT is practically never the arg-max, so every clip decodes the full
224 tokens of whisper.rs:97 -- the worst case for the decoder.  Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import whisper_burn_b200 as wb
from bench import shard_clips, max_over_ranks
from whisper_burn_b200.gguf_file import random_whisper_gguf


def synthetic_mel(cfg, seed):
    rng = np.random.default_rng(seed)
    return np.clip(rng.standard_normal((cfg.n_mels, 2 * cfg.n_audio_ctx)).astype(np.float32) * 0.5, -1.0, 1.5)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--clips", type=int, default=256)
    ap.add_argument("--micro", type=int, default=32, help="clips per transcribe call")
    ap.add_argument("--model", default="large_v3", choices=["large_v3", "medium"])
    ap.add_argument("--language", default="en")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    cfg = wb.WhisperConfig.medium() if a.model == "medium" else wb.WhisperConfig.large_v3()
    blob = random_whisper_gguf(cfg, seed=0)
    model = wb.load_whisper_from_gguf(blob, cfg, dev)
    del blob
    lo, hi = shard_clips(a.clips, rank, world)
    mine = hi - lo
    micro = min(a.micro, max(mine, 1))
    model.reserve(micro)
    host = torch.empty((mine, cfg.n_mels, 2 * cfg.n_audio_ctx), dtype=torch.float32).pin_memory()
    for i in range(mine):
        host[i] = torch.from_numpy(synthetic_mel(cfg, 100 + lo + i))
    # warm-up: one small call (module load, TMA descriptors, plans)
    model.transcribe(host[:min(2, mine)].to(dev), a.language)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    wb.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    n_tok = 0
    first = None
    for s in range(0, mine, micro):
        mel = host[s:s + micro].to(dev, non_blocking=True)          # H2D inside the timed region
        ids, _ = model.transcribe(mel, a.language)                  # token ids come back to the host
        n_tok += sum(len(t) for t in ids)
        if first is None:
            first = ids[0][:8]
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    dt = max_over_ranks(e0.elapsed_time(e1) * 1e-3)
    launches = wb.launch_count()
    tok = torch.tensor([n_tok], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tok)
    if rank == 0:
        print(json.dumps({"metric": "full transcribe 30s-clips/s", "value": a.clips / dt, "unit": "clips/s", "n_gpus": world,
                          "clips": a.clips, "clips_per_gpu": mine, "micro_batch": micro, "seconds": dt, "wall_seconds_rank0": wall,
                          "tokens": int(tok.item()), "tokens_per_s": float(tok.item()) / dt, "scaling": "strong",
                          "gpu_launches_rank0": launches, "first_tokens_clip0": first,
                          "config": {"workload": f"whisper-{a.model} full transcribe (encoder + cross K/V + greedy decode, "
                                                 f"224-token cap), {a.clips} synthetic 30 s clips split by utterance",
                                     "data": "synthetic mel, random-init Q4_0 weights", "timing": "CUDA events, max over ranks, "
                                     "H2D of the mel and D2H of the token ids inside the timed region"},
                          "device_error_flag": int(wb.lib().wq4_device_error_flag(local))}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
