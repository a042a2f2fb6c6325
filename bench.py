#!/usr/bin/env python
"""bench.py - echo waveforms/s of the StofNet inference path (forward + ToF extraction) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Contract (see the task statement): one JSON line on rank 0.  A "step" is one pass of the hot path
over one batch of synthetic waveforms: StofNet.forward (conv stack + semi-global block + sub-pixel
shuffle) followed by mask2coords (NMS + threshold + ToF table).

Workload (config.workload): BASELINE.json configs[1] - StofNet rf_scale_factor=20, 4096 waveforms of
L = 5120 samples per GPU, checkpoint graceful-snow-1553 (the StofNet rf-scale-20 checkpoint; the
literally named wobbly-sponge file is an ESPCN_1D checkpoint), th = 0.015,
nms_win_size = 20.  Rows are independent, so N GPUs each process their own 4096 rows (weak scaling)
and the only collective is the final ToF gather (max-reduce of K_max + all_gather of the table),
issued on a side stream so that it overlaps the next batch's forward (stofnet_b200/sharded.py).

  value     waveforms/s, all ranks, inputs resident in HBM (device timed, max over ranks)
  e2e       same metric through the host-buffer C-ABI call stof_infer_host: pinned host input,
            H2D + forward + peaks + D2H of the ToF table inside the timed region
  roofline  dominant kernel (chain_tc_kernel: conv2..conv12 + conv_last, 65 % of the FLOPs): algorithmic FLOPs
            per launch / mean launch duration from CUDA events recorded on the launch stream inside the timed steps
  parity_check  after the timed loops: 64 rows of the LAST timed step's ToF table against the CPU port of the reference on
            the same rows (index-exact, or margin-classified with the fp64 forward), K_max <= CAP asserted
  cpu_baseline  the oracle's torch-CPU port of the reference path on this box's host cores (bounded row sample;
            rank 0 at N = 1 only)
  other_configs  BASELINE configs 1, 3, 4, 5 (C1 latency, C3 PALA-shaped 16,000 rows per GPU incl. the gather, C4 training
            step at batch 4 and 256, two C5 lengths), each with its own CPU-port figure on a stated subset

--impl reference times that CPU port alone (kind = "port": the reference is a script tree of pure Python/PyTorch that
cannot be installed and is absent on the GPU box; the port re-issues the reference's ATen calls in the reference's order
and is pinned bit-for-bit on the reference's outputs in tests/test_oracle.py), all host threads, 256 rows per step.  That arm
imports nothing of the product: it loads stofnet_b200/synth.py (numpy only) by file path and oracle/torch_port.py.
"""
import argparse
import ctypes as C
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np
import torch

GOLD = os.path.join(ROOT, "tests", "golden")
CKPT = "graceful-snow"
ROWS, L, RF, TH, WIN, R = 4096, 5120, 20, 0.015, 20, 4
CAP = 128                        # initial ToF table width; widened before the timed region if a batch needs more columns
REF_ROWS = 256                   # rows per step of the CPU reference arm (">= 256-row subset")
# algorithmic work per row: MACs = 480,576 L + 163,840 floor(L/80)
FLOP_PER_ROW = 2.0 * (480576 * L + 163840 * (L // 80))
MID_FLOP_PER_POS = 2.0 * 64 * 64 * 7         # one conv2..conv12 layer, per position
CHAIN_FLOP_PER_POS = 11 * MID_FLOP_PER_POS + 2.0 * 64 * 4 * 3     # conv2..conv12 + conv_last (K5 + K6)


def flops_row(length):
    return 2.0 * (480576 * length + 163840 * (length // 80))


def _load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def synth_module():
    """stofnet_b200/synth.py loaded by file path: numpy only, and the reference arm must not import the product package
    (its __init__ maps libstofnet_b200.so)."""
    return _load_by_path("stof_synth", os.path.join(ROOT, "stofnet_b200", "synth.py"))


def peaks_file():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active")
                                                         for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": reasons}


def load_sd(ckpt=CKPT):
    with np.load(os.path.join(GOLD, "weights_%s.npz" % ckpt)) as z:
        return {k: torch.from_numpy(np.ascontiguousarray(z[k])) for k in z.files}


def make_inputs(n_unique, seed, length=L, rf=RF):
    """(n_unique, length) base rows from the seeded, host-independent generator (data: synthetic; 16-bit ADC grid)."""
    return torch.from_numpy(synth_module().synth_waveforms(n_unique, length, seed=seed, rf=rf, q15=True))


def expand_batch(base, rows, shift):
    """rows x L batch in which EVERY row differs: block g of len(base) rows is the base set circularly shifted by
    613 g + shift samples along time and scaled by a per-row gain in [0.8, 1] (detections and K_max are data dependent)."""
    nb, length = base.shape
    out = torch.empty((rows, length), dtype=torch.float32)
    gain = 1.0 - 0.2 * torch.frac(torch.arange(rows, dtype=torch.float64) * 0.6180339887498949).to(torch.float32)
    for g in range((rows + nb - 1) // nb):
        n = min(nb, rows - g * nb)
        out[g * nb:g * nb + n] = torch.roll(base[:n], shifts=(613 * g + shift) % length, dims=1)
    return out * gain.unsqueeze(1)


def reference_batch(rows=REF_ROWS):
    """the first `rows` rows of the first timed batch of the product arm (rank 0)"""
    return expand_batch(make_inputs(256, 3008), rows, 0).unsqueeze(1)


def cpu_port_rate(sd, x, repeats, threads, win=WIN, th=TH):
    """waveforms/s of the oracle's torch-CPU port (forward + mask2coords) on the rows of x (N, 1, L)."""
    from oracle import torch_port
    torch.set_num_threads(threads)
    torch_port.infer(sd, x[:max(1, x.shape[0] // 4)], win, th)       # warm-up
    ts = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        torch_port.infer(sd, x, win, th)
        ts.append(time.perf_counter() - t0)
    return x.shape[0] / float(np.median(ts)), float(np.median(ts))


def run_reference(args, rank, world):
    if rank != 0:
        return
    assert "stofnet_b200" not in sys.modules
    sd = load_sd()
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    from oracle import torch_port
    x = reference_batch()
    for _ in range(max(1, args.warmup)):
        torch_port.infer(sd, x[:64], WIN, TH)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        torch_port.infer(sd, x, WIN, TH)
    dt = time.perf_counter() - t0
    val = REF_ROWS * args.steps / dt
    assert "stofnet_b200" not in sys.modules, "the reference arm must not load the product"
    line = {
        "impl": "reference", "metric": "echo waveforms/sec", "value": val, "unit": "waveforms/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": val, "unit": "waveforms/s", "cores": threads, "kind": "port",
                         "sample": "%d rows of L=%d per step (forward + mask2coords), torch %s CPU, oracle/torch_port.py"
                                   % (REF_ROWS, L, torch.__version__)},
        "e2e": {"value": val, "unit": "waveforms/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def workload_config(gpus):
    return {"workload": "StofNet rf_scale_factor=20 inference (forward + mask2coords), ckpt graceful-snow-1553, "
                        "%d waveforms x L=%d per GPU, th=%g, nms_win_size=%d, upsample_factor=%d"
                        % (ROWS, L, TH, WIN, R),
            "rows_per_gpu": ROWS, "L": L, "rf_scale_factor": RF, "th": TH, "parallelism": "rows sharded x%d" % gpus,
            "l2": "3 distinct input batches rotated (252 MB; every row of a batch differs) and >1 GiB of activations "
                  "per chunk: working set >> 126 MB L2"}


# ----------------------------------------------------------------------------------------------------------------------
# parity check on the timed step's own output
# ----------------------------------------------------------------------------------------------------------------------
def parity_check(sd, x_rows, coords_rows, counts_rows, kmax_step, cap):
    """64 rows of the last timed step: ToF table vs the CPU port of the reference on the same rows."""
    from oracle import torch_port
    pu = _load_by_path("parity_util", os.path.join(ROOT, "tests", "parity_util.py"))
    n = x_rows.shape[0]
    ref_m, ref_co = torch_port.infer(sd, x_rows.unsqueeze(1), WIN, TH)
    ref_idx = torch_port.get_maxima_positions(ref_m, WIN, TH).numpy()
    ours = coords_rows.numpy()
    cnt = counts_rows.numpy().astype(np.int64)
    rows = np.repeat(np.arange(n), np.minimum(cnt, ours.shape[1]))
    pos = np.concatenate([np.rint(ours[i, :min(cnt[i], ours.shape[1])] * R).astype(np.int64) for i in range(n)]) if rows.size else rows
    idx = np.stack([rows, pos], axis=1) if rows.size else np.zeros((0, 2), np.int64)
    n_diff, n_marg, _ = pu.classify_mismatches(sd, x_rows.numpy(), idx, ref_idx, TH, ref_masks32=ref_m.squeeze(1).numpy())
    table_equal = False
    if n_diff == 0 and ref_co.dim() == 2:
        k = ref_co.shape[1]
        table_equal = bool(np.array_equal(ours[:, :k], ref_co.numpy()) and not ours[:, k:].any())
    return {"rows": int(n), "detections": int(ref_idx.shape[0]), "mismatches": int(n_diff - n_marg),
            "boundary_flips": int(n_marg), "tof_table_identical": table_equal, "kmax": int(kmax_step), "cap": int(cap),
            "checker": "oracle/torch_port.py on this box's CPU, margins by its fp64 re-run (tests/parity_util.py)"}


# ----------------------------------------------------------------------------------------------------------------------
# BASELINE configs 1, 3, 4, 5 (evidence lines beside the headline; each failure is reported, not raised)
# ----------------------------------------------------------------------------------------------------------------------
def other_configs(sb, dev, rank, world, cpu):
    import torch.distributed as dist
    from stofnet_b200._lib import lib, check
    from stofnet_b200.engine import stream_ptr
    from stofnet_b200.sharded import TofGatherPipeline
    syn = synth_module()
    threads = os.cpu_count() or 1
    out = {}

    def dev_ms(fn, warm, reps):
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
        t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def infer_step(model, x, th, cap, pipe=None):
        n, length = x.shape[0], x.shape[-1]
        coords = [torch.empty((n, cap), dtype=torch.float32, device=dev) for _ in range(2)]
        counts = torch.empty((n,), dtype=torch.int32, device=dev)
        stats = torch.zeros((2,), dtype=torch.int64, device=dev)
        state = {"i": 0, "pending": None}

        def step():
            c = coords[state["i"] % 2]
            state["i"] += 1
            with torch.no_grad():
                mask = model(x)
            stats.zero_()
            check(lib.stof_peaks(mask.data_ptr(), n, R * length, WIN, 1 if th is None else 0, 0.0 if th is None else th, R,
                                 c.data_ptr(), None, cap, counts.data_ptr(), stats.data_ptr(), stream_ptr(dev)))
            if pipe is not None:
                slot = pipe.submit(c, stats[:1])
                if state["pending"] is not None:
                    pipe.collect(*state["pending"])
                state["pending"] = (slot, c)

        def drain():
            if pipe is not None and state["pending"] is not None:
                pipe.collect(*state["pending"])
                state["pending"] = None
        return step, drain

    def guarded(name, fn):
        try:
            out[name] = fn()
        except Exception as e:            # evidence lines must not take the headline down
            out[name] = {"error": "%s: %s" % (type(e).__name__, e)}
        torch.cuda.empty_cache()

    arm = sb.build_model(load_sd("different-armadillo"), dev)
    sd_arm = load_sd("different-armadillo")

    # ---- C1: N = 1, L = 2560 (rf 10), different-armadillo, th = None; latency of one eval step from host buffers
    def c1():
        x = torch.from_numpy(syn.synth_waveforms(1, 2560, seed=3008, rf=10, q15=True)).pin_memory()
        for _ in range(5):
            arm.infer_host(x, WIN, None, cap=16)
        ts = []
        for _ in range(30):
            t0 = time.perf_counter()
            arm.infer_host(x, WIN, None, cap=16)
            ts.append(time.perf_counter() - t0)
        d = {"config": "C1: batch_size=1, L=2560, ckpt different-armadillo, th=None, host buffers in / ToF table out (stof_infer_host)",
             "latency_ms": float(np.median(ts)) * 1e3, "waveforms_per_s": 1.0 / float(np.median(ts))}
        if cpu:
            rate, med = cpu_port_rate(sd_arm, x.unsqueeze(1), 7, threads, th=None)
            d["cpu_port"] = {"latency_ms": med * 1e3, "waveforms_per_s": rate, "cores": threads, "sample": "the same row, median of 7"}
        return d
    if world == 1:
        guarded("C1", c1)

    # ---- C2 with the checkpoint BASELINE.json names literally (wobbly-sponge is an ESPCN_1D checkpoint): 4096 rows x
    #      L = 5120 per GPU, th = 0.2, forward (espcn_tc.cu: conv2 / conv3 on tcgen05) + mask2coords, device-resident
    def c2_espcn():
        from oracle import torch_port
        sd_e = load_sd("wobbly-sponge")
        model = sb.build_model(sd_e, dev)
        x = expand_batch(make_inputs(256, 3008 + rank), ROWS, 0).unsqueeze(1).contiguous().to(dev)
        step, _ = infer_step(model, x, 0.2, 256)
        ms = dev_ms(lambda: step(), 2, 5)
        d = {"config": "C2 with ckpt wobbly-sponge (ESPCN_1D): %d rows x L=%d per GPU, th=0.2, forward + mask2coords" % (ROWS, L),
             "ms_per_step": ms, "waveforms_per_s": ROWS * world / (ms / 1e3)}
        if cpu:
            xs = expand_batch(make_inputs(256, 3008), 256, 0).unsqueeze(1)
            torch.set_num_threads(threads)
            torch_port.infer(sd_e, xs[:64], WIN, 0.2, espcn=True)
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                torch_port.infer(sd_e, xs, WIN, 0.2, espcn=True)
                ts.append(time.perf_counter() - t0)
            med = float(np.median(ts))
            d["cpu_port"] = {"waveforms_per_s": 256 / med, "cores": threads, "sample": "256 rows, median of 3 (%.2f s each)" % med}
        del x
        model._engine.ws = None
        return d
    guarded("C2_espcn", c2_espcn)

    # ---- C3: PALA-shaped frames (128 channels folded into rows), 16,000 rows per GPU, th = None, incl. the ToF gather
    def c3():
        rows = 16000
        base = torch.from_numpy(syn.pala_shaped(128, 8, 2560, seed=3008 + rank, rf=10, q15=True))     # 8 frames = 1024 rows
        x = expand_batch(base, rows, 0).unsqueeze(1).contiguous().to(dev)
        pipe = TofGatherPipeline(rows, 64, dev) if world > 1 else None
        step, drain = infer_step(arm, x, None, 64, pipe)
        ms = dev_ms(lambda: step(), 1, 3)
        drain()
        d = {"config": "C3: PALA-shaped, 128 channels x %d frames = %d rows x L=2560 (16,000 rows per GPU), ckpt different-armadillo, "
                       "th=None, forward + mask2coords + ToF gather" % (125 * world, rows * world),
             "ms_per_step": ms, "waveforms_per_s": rows * world / (ms / 1e3),
             "algorithmic_tflops": flops_row(2560) * rows * world / (ms / 1e3) / 1e12}
        if pipe is not None:
            d["gather"] = {"stride_columns": pipe.k_opt, "fallbacks": pipe.fallbacks}
        if cpu:
            rate, med = cpu_port_rate(sd_arm, expand_batch(base, 256, 0).unsqueeze(1), 3, threads, th=None)
            d["cpu_port"] = {"waveforms_per_s": rate, "cores": threads, "sample": "256 of the rows, median of 3 (%.2f s each)" % med}
        del x
        arm._engine.ws = None
        return d
    guarded("C3", c3)

    # ---- C4: training step (forward + loss + backward + AdamW [+ gradient all-reduce]), rf 10, batch 4 and 256 per GPU
    def c4():
        res = {"config": "C4: training step forward + loss + backward + AdamW%s, ckpt different-armadillo, L=2560, synthetic chirp "
                         "batches" % (" + gradient all-reduce" if world > 1 else "")}
        for rows in (4, 256):
            model = sb.build_model(load_sd("different-armadillo"), dev)
            model.train()
            tr = sb.Trainer(model, lr=5e-4, weight_decay=1e-8, lambda_value=1e-2)
            x = torch.from_numpy(syn.synth_waveforms(rows, 2560, seed=11 + rank, rf=10, q15=True)).unsqueeze(1).to(dev)
            rng = np.random.default_rng(5 + rank)
            gt = torch.from_numpy(rng.integers(100, 4 * 2560 - 100, size=(rows, 1, 3))).to(dev)
            ms = dev_ms(lambda: tr.step(x, gt), 2, 5)
            # forward + data gradient + weight gradient of every conv: 3 x the forward's algorithmic FLOPs
            res["batch_%d" % rows] = {"ms_per_step": ms, "rows_per_s": rows * world / (ms / 1e3),
                                      "algorithmic_tflops": 3 * flops_row(2560) * rows * world / (ms / 1e3) / 1e12}
            if world == 1:
                # the same step replayed from ONE CUDA graph (Trainer.capture): what is left when the ~100 launches cost no host time
                tr.capture(x, gt)
                ms = dev_ms(lambda: tr.step_graph(x, gt), 3, 10)
                res["batch_%d" % rows]["cuda_graph"] = {"ms_per_step": ms, "rows_per_s": rows / (ms / 1e3)}
            del tr, model
        if cpu:
            res["cpu_port"] = cpu_train_rate(sd_arm, syn, threads)
        return res
    guarded("C4", c4)

    # ---- C5: two ends of the length sweep (raw 256 x rf 10 and raw 8192 x rf 20)
    def c5():
        res = {"config": "C5: forward + mask2coords at the ends of the length sweep, ~16.8 M positions per GPU and step"}
        snow = sb.build_model(load_sd("graceful-snow"), dev)
        for length, rf, model, ck, th in ((2560, 10, arm, "different-armadillo", None), (163840, 20, snow, "graceful-snow", TH)):
            rows = max(64, (16 << 20) // length)
            base = torch.from_numpy(syn.synth_waveforms(min(rows, 128), length, seed=3008 + rank, rf=rf, q15=True))
            x = expand_batch(base, rows, 0).unsqueeze(1).contiguous().to(dev)
            step, _ = infer_step(model, x, th, 2048 if length > 100000 else 128)
            ms = dev_ms(lambda: step(), 1, 3)
            d = {"rows_per_gpu": rows, "ckpt": ck, "th": th, "ms_per_step": ms, "waveforms_per_s": rows * world / (ms / 1e3),
                 "algorithmic_tflops": flops_row(length) * rows * world / (ms / 1e3) / 1e12}
            if cpu:
                n_cpu = 64 if length == 2560 else 1
                rate, med = cpu_port_rate(load_sd(ck), expand_batch(base, n_cpu, 0).unsqueeze(1), 3, threads, th=th)
                d["cpu_port"] = {"waveforms_per_s": rate, "cores": threads, "sample": "%d row(s), median of 3 (%.2f s each)" % (n_cpu, med)}
            res["L_%d" % length] = d
            del x
            model._engine.ws = None
            torch.cuda.empty_cache()
        return res
    guarded("C5", c5)
    return out


def cpu_train_rate(sd, syn, threads):
    """The reference's training step (main.py:217-248) through the port + torch autograd + torch.optim.AdamW on the host."""
    import torch.nn.functional as F
    from oracle import torch_port
    torch.set_num_threads(threads)
    rows = 4
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    opt = torch.optim.AdamW(list(params.values()), lr=5e-4, weight_decay=1e-8)
    x = torch.from_numpy(syn.synth_waveforms(rows, 2560, seed=11, rf=10, q15=True)).unsqueeze(1)
    gt = torch.from_numpy(np.random.default_rng(5).integers(100, 4 * 2560 - 100, size=(rows, 1, 3)))
    g = torch.tensor([0.00443185, 0.05399097, 0.24197072, 0.39894228, 0.24197072, 0.05399097, 0.00443185]).view(1, 1, -1)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        opt.zero_grad()
        pred = torch_port.stofnet_forward(params, x)
        mask = torch.zeros_like(pred)
        mask.scatter_(2, gt, 1)
        blur = F.conv1d(mask, g, padding=3)
        blur = blur / blur.max() * 20
        loss = F.mse_loss(pred.squeeze(1), blur.squeeze(1)) + 1e-2 * pred.abs().mean()
        loss.backward()
        opt.step()
        ts.append(time.perf_counter() - t0)
    med = float(np.median(ts[1:]))
    return {"batch_4": {"ms_per_step": med * 1e3, "rows_per_s": rows / med}, "cores": threads,
            "sample": "batch 4 (config.yaml:11), median of 3 steps; torch autograd + torch.optim.AdamW through oracle/torch_port.py"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true", help="skip BASELINE configs 1, 3, 4, 5 (other_configs)")
    ap.add_argument("--layers-per-launch", type=int, default=0, help="conv layers fused per chain launch (0 = library default)")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp16"],
                    help="fp32 (default, the headline: index-exact) or fp16 (one-pass reduced-precision mode, stated tolerance)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch.distributed as dist
    import stofnet_b200 as sb
    from stofnet_b200._lib import lib, check, PROF_CLASSES, PROF_NAMES
    from stofnet_b200.engine import stream_ptr
    from stofnet_b200.sharded import TofGatherPipeline

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU path)"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        # rank 0's stdout carries ONE JSON line: keep NCCL's version banner (NCCL_DEBUG=VERSION on some boxes) off it
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
            os.environ.pop("NCCL_DEBUG")
        dist.init_process_group("nccl", device_id=dev)

    sd = load_sd()
    model = sb.build_model(sd, dev)
    model.set_precision(args.precision)
    eng = model._engine
    if args.layers_per_launch:
        eng.set_path(eng.path, args.layers_per_launch)          # tuning knob: conv layers fused per chain launch
    base = make_inputs(256, 3008 + rank)
    batches_host = [expand_batch(base, ROWS, s).pin_memory() for s in (0, 7, 13)]
    batches_dev = [b.to(dev).unsqueeze(1).contiguous() for b in batches_host]

    # ToF table width: every distinct batch is run once (untimed) and the table widened until no batch is truncated
    # (stof_peaks reports the true K_max even when it exceeds the width it was given)
    cap = CAP
    probe_c = torch.empty((ROWS, cap), dtype=torch.float32, device=dev)
    probe_n = torch.empty((ROWS,), dtype=torch.int32, device=dev)
    probe_s = torch.zeros((2,), dtype=torch.int64, device=dev)
    with torch.no_grad():
        for x in batches_dev:
            mask = eng.forward(model, x)
            probe_s.zero_()
            check(lib.stof_peaks(mask.data_ptr(), ROWS, R * L, WIN, 0, TH, R, probe_c.data_ptr(), None, cap,
                                 probe_n.data_ptr(), probe_s.data_ptr(), stream_ptr(dev)))
            cap = max(cap, (int(probe_s[0].item()) + 31) // 32 * 32)
        del mask
    kc = torch.tensor([cap], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(kc, op=dist.ReduceOp.MAX)
    cap = int(kc.item())
    del probe_c, probe_n, probe_s
    coords = [torch.empty((ROWS, cap), dtype=torch.float32, device=dev) for _ in range(2)]
    counts = [torch.empty((ROWS,), dtype=torch.int32, device=dev) for _ in range(2)]
    stats = [torch.zeros((2,), dtype=torch.int64, device=dev) for _ in range(2)]
    # final ToF gather (north_star): global K_max + fixed-stride table, on a side stream behind the next batch's forward
    pipe = TofGatherPipeline(ROWS, cap, dev) if world > 1 else None
    pending = [None]

    def step(i):
        x = batches_dev[i % len(batches_dev)]
        b = i % 2
        mask = eng.forward(model, x)
        stats[b].zero_()
        check(lib.stof_peaks(mask.data_ptr(), ROWS, R * L, WIN, 0, TH, R, coords[b].data_ptr(), None, cap,
                             counts[b].data_ptr(), stats[b].data_ptr(), stream_ptr(dev)))
        if pipe is not None:
            slot = pipe.submit(coords[b], stats[b][:1])
            if pending[0] is not None:
                pipe.collect(*pending[0])             # the PREVIOUS step's global table (its exchange ran behind this forward)
            pending[0] = (slot, coords[b])
        return b

    def drain():
        if pipe is not None and pending[0] is not None:
            pipe.collect(*pending[0])
            pending[0] = None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        for i in range(args.warmup):
            step(i)
        drain()
        barrier()
        check(lib.stof_profile_enable(eng.handle, 1))
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        l0 = lib.stof_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        last_b = 0
        for i in range(args.steps):
            last_b = step(args.warmup + i)
        drain()                                        # the last step's gather is inside the timed region
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = lib.stof_launch_count() - l0
        prof_ms = (C.c_double * PROF_CLASSES)()
        prof_n = (C.c_int64 * PROF_CLASSES)()
        check(lib.stof_profile_collect(eng.handle, prof_ms, prof_n))
        check(lib.stof_profile_enable(eng.handle, 0))
        last_i = args.warmup + args.steps - 1
        kmax_step = int(stats[last_b][0].item())
        assert kmax_step <= cap, "K_max=%d exceeds cap=%d: the timed step's ToF table is truncated" % (kmax_step, cap)
        dev_coords = coords[last_b].cpu()
        dev_counts = counts[last_b].cpu()

        # ---- e2e: host buffers through stof_infer_host (H2D + forward + peaks + D2H inside the timing)
        coords_h = torch.empty((ROWS, cap), dtype=torch.float32).pin_memory()
        counts_h = torch.empty((ROWS,), dtype=torch.int32).pin_memory()
        kmax = C.c_int(0)

        def e2e_step(i):
            xh = batches_host[i % len(batches_host)]
            check(lib.stof_infer_host(eng.handle, xh.data_ptr(), ROWS, L, WIN, 0, TH, coords_h.data_ptr(), cap,
                                      counts_h.data_ptr(), None, C.byref(kmax)))

        for i in range(max(1, args.warmup - 1)):
            e2e_step(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            e2e_step(i)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        clocks = sampler.stop() if rank == 0 else None
        assert kmax.value <= cap, "K_max=%d exceeds cap=%d in stof_infer_host" % (kmax.value, cap)
        # the host-buffer call and the device-resident path must give the same table on the same batch
        e2e_step(last_i)
        e2e_same = bool(torch.equal(coords_h, dev_coords) and torch.equal(counts_h, dev_counts))

    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    value = ROWS * world * args.steps / (ms / 1e3)
    e2e_val = ROWS * world * args.steps / (e2e_ms / 1e3)

    line = None
    if rank == 0:
        peaks, peak_src = peaks_file()
        # dominant kernel = chain_tc_kernel: the fused 64->64 k7 stack (conv2..conv12) + conv_last/shuffle, launched
        # per segment of layers (profile classes conv_mid = segments of mid layers, conv_last = the tail segment)
        ch_n = int(prof_n[4]) + int(prof_n[5])
        ch_ms = (float(prof_ms[4]) + float(prof_ms[5])) / max(1, ch_n)
        chain_flops_step = CHAIN_FLOP_PER_POS * ROWS * L                    # algorithmic, per step
        flops_per_launch = chain_flops_step * args.steps / max(1, ch_n)
        achieved_tf = flops_per_launch / (ch_ms / 1e3) / 1e12 if ch_n else 0.0
        peak_tf = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops")))
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "chain_traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj = json.load(f)
            # ncu dram bytes (read + write) per position summed over the chain launches of one forward, averaged per launch
            traffic = float(tj["dram_bytes_per_position_all_segments"]) * ROWS * L * args.steps / max(1, ch_n)
        share = {PROF_NAMES[c]: round(float(prof_ms[c]) / args.steps, 3) for c in range(PROF_CLASSES) if prof_n[c]}
        # MMA FLOPs issued per algorithmic FLOP: 3 for the fp32-faithful hi/lo split, 1 for the one-pass fp16 mode
        issue_x = 1.0 if args.precision == "fp16" else 3.0
        line = {
            "metric": "echo waveforms/sec", "value": value, "unit": "waveforms/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16 (one pass, fp32 accumulate; opt-in tolerance mode)" if args.precision == "fp16" else "f32",
            "data": "synthetic", "config": workload_config(world),
            "e2e": {"value": e2e_val, "unit": "waveforms/s", "h2d_bytes_per_step": ROWS * L * 4,
                    "d2h_bytes_per_step": ROWS * cap * 4 + ROWS * 4 + 16, "ms_per_step": e2e_ms / args.steps,
                    "same_table_as_device_path": e2e_same},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor",
                         "kernel": "chain_tc_kernel (conv2..conv12 64->64 k7 + conv_last/shuffle; tcgen05 kind::f16, %s)"
                                   % ("one pass on fp16-rounded operands" if issue_x == 1.0 else
                                      "fp32 operands as fp16 hi/lo pairs: 3 issued MMA FLOPs per algorithmic FLOP"),
                         "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved_tf / peak_tf if peak_tf else None, "traffic": traffic,
                         "issued_tflops": issue_x * achieved_tf, "issued_frac": issue_x * achieved_tf / peak_tf if peak_tf else None,
                         "peak_source": "%s: dense 16-bit tensor peak, sustained figure of MEASURED_PEAKS.json (kernel timed "
                                        "inside a long step); 'achieved' counts algorithmic FLOPs, 'issued' the fp16 MMAs "
                                        "the exact-fp32 split needs" % peak_src,
                         "launch_ms": ch_ms, "launches": ch_n,
                         "algorithmic_flops_per_launch": flops_per_launch},
            "kernel_ms_per_step": share,
            "algorithmic_tflops": FLOP_PER_ROW * value / 1e12,
        }
        if pipe is not None:
            line["tof_gather"] = {"stride_columns": pipe.k_opt, "cap": cap, "fallbacks": pipe.fallbacks,
                                  "overlapped": "side stream, behind the next batch's forward"}
        if args.precision == "fp32":
            x64 = batches_host[last_i % len(batches_host)][:64]
            line["parity_check"] = parity_check(sd, x64, dev_coords[:64], dev_counts[:64], kmax_step, cap)
            assert line["parity_check"]["mismatches"] == 0, "bench parity check failed: %r" % (line["parity_check"],)
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            rate, med = cpu_port_rate(sd, reference_batch(), 3, threads)
            line["cpu_baseline"] = {"value": rate, "unit": "waveforms/s", "cores": threads, "kind": "port",
                                    "sample": "%d rows of L=%d, forward + mask2coords, median of 3 (%.2f s each), "
                                              "torch %s CPU via oracle/torch_port.py" % (REF_ROWS, L, med, torch.__version__)}
    if not args.no_other_configs:
        del batches_dev
        eng.ws = None
        torch.cuda.empty_cache()
        oc = other_configs(sb, dev, rank, world, cpu=(rank == 0 and world == 1 and not args.no_cpu_baseline))
        if line is not None:
            line["other_configs"] = oc
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
