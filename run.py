#!/usr/bin/env python
"""run.py - minimal evaluation runner mirroring the reference's eval call sites (main.py:291-347).

    python run.py [config.yaml] key=value ...      e.g.  python run.py model_file=graceful-snow rf_scale_factor=20 th=0.015

    cfg            config.yaml + key=value overrides                                   main.py:31-34
    checkpoint     first file in cfg.ckpt_dir starting with model_file.split('_')[0]   main.py:174-177
    frame          (batch_size, raw_len) synthetic RF rows -> upscale_1d(rf_scale_factor)   datasets/chirp_dataset.py:113
    masks_pred + es_sample   model.infer_host(frame, nms_win_size, th)                 main.py:307,314,320
    toa_errs       toa_rmse(gt_sample, es_sample, tol=cfg.etol)                        main.py:347

The reference's datasets are absent, so frames are the synthetic chirp-like echoes of stofnet_b200.synth with their echo
centres as ground truth.  Everything numerical runs in libstofnet_b200.so on the GPU; there is no CPU path.
Prints one JSON line with the reference's metric names (RMSE / precision / recall / Jaccard averaged over the rows).
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def evaluate(cfg, raw_len=256, batches=2):
    import stofnet_b200 as sb
    from stofnet_b200.synth import synth_waveforms

    if not torch.cuda.is_available():
        raise RuntimeError("run.py needs a CUDA device (stofnet_b200 has no CPU path)")
    dev = torch.device(cfg.device if cfg.device and str(cfg.device).startswith("cuda") else "cuda")
    if cfg.model_file is None:
        raise ValueError("model_file=<checkpoint prefix> is required (main.py:173-177 loads it the same way)")
    ckpt = sb.checkpoint.find_checkpoint(cfg.ckpt_dir, str(cfg.model_file))
    model = sb.build_model(sb.load_checkpoint(ckpt), dev)
    rf, r = cfg.rf_scale_factor, cfg.upsample_factor
    if r != model.upsample_factor:
        raise ValueError("upsample_factor=%r but the checkpoint's conv_last has %d output channels" % (r, model.upsample_factor))
    rows = []
    for b in range(batches):
        raw, centres = synth_waveforms(int(cfg.batch_size), raw_len, seed=int(cfg.seed) + b, rf=1, return_centres=True)
        # dataset side (datasets/chirp_dataset.py:113,120-122): interpolate the raw row by rf_scale_factor, scale the GT alike
        frame = sb.upscale_1d(torch.from_numpy(raw).to(dev), rf)
        gt_sample = torch.from_numpy(centres).to(dev) * rf
        gt_sample[(gt_sample <= 0) | torch.isnan(gt_sample)] = 0                    # main.py:309
        es_sample, counts, kmax, _ = model.infer_host(frame.cpu().pin_memory(), cfg.nms_win_size, cfg.th)
        es = es_sample[:, :max(kmax, 1)].to(dev)
        rows.append(sb.toa_rmse(gt_sample, es, tol=cfg.etol).cpu())                   # main.py:347
    t = torch.cat(rows)
    return {"checkpoint": os.path.basename(ckpt), "model": type(model).__name__, "rows": int(t.shape[0]),
            "L": int(raw_len * rf), "rf_scale_factor": rf, "th": cfg.th, "etol": cfg.etol,
            "rmse": float(np.nanmean(t[:, 0].numpy())), "precision": float(np.nanmean(t[:, 1].numpy())),
            "recall": float(np.nanmean(t[:, 2].numpy())), "jaccard": float(np.nanmean(t[:, 3].numpy())),
            "tp": int(t[:, 4].sum()), "fp": int(t[:, 5].sum()), "fn": int(t[:, 6].sum())}


def main(argv):
    import stofnet_b200 as sb
    path = os.path.join(ROOT, "config.yaml")
    if argv and "=" not in argv[0]:
        path, argv = argv[0], argv[1:]
    cfg = sb.load_config(path, argv)
    print(json.dumps(evaluate(cfg)))


if __name__ == "__main__":
    main(sys.argv[1:])
