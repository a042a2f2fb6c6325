#!/usr/bin/env python
"""Throughput of the path over BASELINE.json's other configs (parity-test / evidence lines, not the bench line):

  C5  waveform length sweep: raw length 256..8192 x rf_scale_factor 10 / 20  (L = 2,560 .. 163,840)
  C3  synthetic PALA-shaped batch, 128 channels x 1000 frames = 128,000 rows, L = 2560 (rf 10), rows sharded over ranks
  C4  training step (forward + loss + backward + AdamW), rf 10, synthetic chirp batches, gradient all-reduce over ranks

One JSON line per case on rank 0 (device-timed with CUDA events, max over ranks; inputs resident in HBM).
    python tools/sweep.py [c5] [c3] [c4]            (or under torchrun for N > 1)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

import stofnet_b200 as sb
from stofnet_b200._lib import lib, check
from stofnet_b200.engine import stream_ptr
from stofnet_b200.synth import synth_waveforms, pala_shaped

GOLD = os.path.join(ROOT, "tests", "golden")
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def model_for(rf):
    name = "different-armadillo" if rf == 10 else "graceful-snow"
    th = None if rf == 10 else 0.015
    return sb.build_model(sb.load_checkpoint(os.path.join(GOLD, "weights_%s.npz" % name)), dev), name, th


def timed(fn, warm=2, reps=3):
    for _ in range(warm):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def emit(d):
    if rank == 0:
        d["n_gpus"] = world
        print(json.dumps(d), flush=True)


def flops_row(L):
    return 2.0 * (480576 * L + 163840 * (L // 80))


def infer_case(model, x, th, cap=256):
    N, L = x.shape[0], x.shape[-1]
    coords = torch.empty((N, cap), dtype=torch.float32, device=dev)
    counts = torch.empty((N,), dtype=torch.int32, device=dev)
    stats = torch.zeros((2,), dtype=torch.int64, device=dev)
    none = 1 if th is None else 0

    def step():
        with torch.no_grad():
            mask = model(x)
        stats.zero_()
        check(lib.stof_peaks(mask.data_ptr(), N, 4 * L, 20, none, 0.0 if th is None else th, 4, coords.data_ptr(), None, cap,
                             counts.data_ptr(), stats.data_ptr(), stream_ptr(dev)))
        if world > 1:
            k = stats[:1].clone()
            dist.all_reduce(k, op=dist.ReduceOp.MAX)
    return step


which = set(sys.argv[1:]) or {"c5", "c3", "c4"}

if "c5" in which:
    for rf in (10, 20):
        model, name, th = model_for(rf)
        for raw in (256, 512, 1024, 2048, 4096, 8192):
            L = raw * rf
            rows = max(64, min(8192, (16 << 20) // L))          # ~16 Mi positions per rank and step
            x = torch.from_numpy(synth_waveforms(min(rows, 128), L, seed=3008 + rank, rf=rf))
            x = x.repeat((rows + x.shape[0] - 1) // x.shape[0], 1)[:rows].unsqueeze(1).contiguous().to(dev)
            ms = timed(infer_case(model, x, th))
            emit({"case": "C5", "rf_scale_factor": rf, "raw_length": raw, "L": L, "rows_per_gpu": rows, "ckpt": name,
                  "ms_per_step": ms, "waveforms_per_s": rows * world / (ms / 1e3),
                  "algorithmic_tflops": flops_row(L) * rows * world / (ms / 1e3) / 1e12})
            del x
            model._engine.ws = None
            torch.cuda.empty_cache()

if "c3" in which:
    model, name, _ = model_for(10)
    frames_total, channels, L = 1000, 128, 2560
    from stofnet_b200.sharded import shard_rows
    lo, hi = shard_rows(frames_total, world, rank)
    base = torch.from_numpy(pala_shaped(channels, 8, L, seed=3008 + rank, rf=10))           # 8 distinct frames, tiled
    n_rows = (hi - lo) * channels
    x = base.repeat((n_rows + base.shape[0] - 1) // base.shape[0], 1)[:n_rows].unsqueeze(1).contiguous().to(dev)
    for th in (None, 0.5):
        ms = timed(infer_case(model, x, th), warm=1, reps=2)
        emit({"case": "C3", "rows_total": frames_total * channels, "rows_per_gpu": n_rows, "L": L, "th": th, "ckpt": name,
              "ms_per_step": ms, "waveforms_per_s": frames_total * channels / (ms / 1e3),
              "algorithmic_tflops": flops_row(L) * frames_total * channels / (ms / 1e3) / 1e12})
    del x
    model._engine.ws = None
    torch.cuda.empty_cache()

if "c4" in which:
    for rows in (4, 64, 256):
        model, name, _ = model_for(10)
        model.train()
        L = 2560
        tr = sb.Trainer(model, lr=5e-4, weight_decay=1e-8, lambda_value=1e-2)
        x = torch.from_numpy(synth_waveforms(rows, L, seed=11 + rank, rf=10)).unsqueeze(1).to(dev)
        rng = np.random.default_rng(5 + rank)
        gt = torch.from_numpy(rng.integers(100, 4 * L - 100, size=(rows, 1, 3))).to(dev)
        ms = timed(lambda: tr.step(x, gt), warm=2, reps=5)
        emit({"case": "C4", "rows_per_gpu": rows, "L": L, "ckpt": name, "ms_per_step": ms,
              "waveforms_per_s": rows * world / (ms / 1e3), "step": "forward+loss+backward+AdamW" + ("+allreduce" if world > 1 else "")})

if world > 1:
    dist.destroy_process_group()
