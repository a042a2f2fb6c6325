"""Host logic of the multi-GPU path (stofnet_b200/sharded.py) on CPU: world_size-2 and -3 gloo groups.

The per-rank compute is CUDA-only, so here each rank's local ToF table comes from the oracle applied to
its shard of a reference-generated mask (tests may use the oracle as the checker / stand-in); what is under
test is the partitioning and the global gather: K_max max-reduce, fixed-stride all_gather, ragged shards,
and the reference's (N, 1, 1) result when no rank detects anything (utils/mask2samples.py:87-88)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN, load_npz

from stofnet_b200.sharded import gather_tof_table, shard_rows


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _masks(n_rows):
    """(n_rows, 1, M) reference masks: golden rows tiled with per-row circular shifts so that rows differ
    and per-row detection counts vary."""
    g = load_npz("forward_graceful-snow.npz")
    base = g["c2_mask"]                                   # (2, 10240) fp32 from the reference
    rows = [np.roll(base[i % base.shape[0]], 37 * i) * (1.0 + 0.01 * i) for i in range(n_rows)]
    return np.stack(rows).astype(np.float32)[:, None, :]


def _worker(rank, world, port, n_rows, th, empty, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import c_oracle
        masks = _masks(n_rows)
        if empty:
            masks = np.zeros_like(masks)
        lo, hi = shard_rows(n_rows, world, rank)
        if hi > lo:
            local = c_oracle.mask2coords(masks[lo:hi], 20, th, 4)
        else:
            local = np.zeros((0, 1), np.float32)
        if local.ndim == 3:                               # (n, 1, 1): nothing detected on this rank
            local_t, kmax = torch.zeros((hi - lo, 1)), 0
        else:
            local_t, kmax = torch.from_numpy(np.ascontiguousarray(local)), local.shape[1]
        # a rank-dependent capacity (as stof_peaks callers have) must not matter
        pad = torch.zeros((hi - lo, local_t.shape[1] + 3 * rank))
        pad[:, :local_t.shape[1]] = local_t
        table = gather_tof_table(pad, kmax, n_rows)
        np.save(os.path.join(out_dir, "table_%d.npy" % rank), table.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_rows,th,empty", [
    (2, 8, 0.5, False),        # equal shards -> all_gather_into_tensor
    (2, 7, 0.015, False),      # ragged shards
    (3, 8, None, False),       # row-max mode, K_max = 1 typically
    (2, 6, 0.5, True),         # nothing detected anywhere -> zeros((N, 1, 1))
    (3, 2, 0.5, False),        # more ranks than rows on one rank (empty shard)
])
def test_gather_matches_undivided_batch(tmp_path, world, n_rows, th, empty):
    from oracle import c_oracle
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_rows, th, empty, str(tmp_path)), nprocs=world, join=True)
    masks = _masks(n_rows)
    if empty:
        masks = np.zeros_like(masks)
    want = c_oracle.mask2coords(masks, 20, th, 4)
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), "table_%d.npy" % r))
        assert got.shape == want.shape, (got.shape, want.shape)
        assert np.array_equal(got, want)


def _pipeline_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import c_oracle
        from stofnet_b200.sharded import TofGatherPipeline
        n_rows, cap = 8, 96
        lo, hi = shard_rows(n_rows, world, rank)
        pipe = TofGatherPipeline(hi - lo, cap, torch.device("cpu"))
        # a stream of batches: the stride narrows to the first batches' K_max, then a batch grows past the optimistic stride (fallback);
        # the last batch detects nothing anywhere
        plan = [(0.5, 1.0), (0.5, 1.0), (0.5, 1.0), (0.015, 1.0), (0.5, 0.0)]
        pending = None
        for step in range(len(plan) + 1):
            if step < len(plan):
                th, gain = plan[step]
                masks = _masks(n_rows) * gain
                local = c_oracle.mask2coords(masks[lo:hi], 20, th, 4)
                tab = torch.zeros((hi - lo, cap))
                kmax = 0
                if local.ndim == 2:
                    kmax = local.shape[1]
                    tab[:, :kmax] = torch.from_numpy(np.ascontiguousarray(local))
                slot = pipe.submit(tab, torch.tensor([kmax]))
            if pending is not None:                     # collect the PREVIOUS step, one step late (the overlapped schedule)
                pstep, pslot, ptab = pending
                out = pipe.collect(pslot, ptab)
                np.save(os.path.join(out_dir, "pipe_%d_%d.npy" % (rank, pstep)), out.numpy())
            pending = (step, slot, tab) if step < len(plan) else None
        np.save(os.path.join(out_dir, "pipe_fallbacks_%d.npy" % rank), np.array([pipe.fallbacks, pipe.k_opt]))
    finally:
        dist.destroy_process_group()


def test_overlapped_gather_pipeline_matches_undivided_batch(tmp_path):
    """TofGatherPipeline (side-stream gather with the previous steps' K_max as the optimistic stride) on a world-size-2
    gloo group: every step's table equals mask2coords on the undivided batch, including the fallback when a batch needs
    more columns than the stride, and the (N, 1, 1) result when nothing is detected."""
    from oracle import c_oracle
    world, n_rows = 2, 8
    mp.spawn(_pipeline_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    plan = [(0.5, 1.0), (0.5, 1.0), (0.5, 1.0), (0.015, 1.0), (0.5, 0.0)]
    for step, (th, gain) in enumerate(plan):
        want = c_oracle.mask2coords(_masks(n_rows) * gain, 20, th, 4)
        for r in range(world):
            got = np.load(os.path.join(str(tmp_path), "pipe_%d_%d.npy" % (r, step)))
            assert got.shape == want.shape, (step, got.shape, want.shape)
            assert np.array_equal(got, want), step
    fb = np.load(os.path.join(str(tmp_path), "pipe_fallbacks_0.npy"))
    assert fb[0] >= 1                                    # step 3 needed more columns than the narrowed stride


def test_shard_rows_partition():
    for n in (0, 1, 7, 128000, 4096):
        for w in (1, 2, 3, 8):
            blocks = [shard_rows(n, w, r) for r in range(w)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for a, b in zip(blocks, blocks[1:]):
                assert a[1] == b[0]
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_rows(4, 2, 2)


def test_single_process_path():
    t = torch.tensor([[1.0, 2.5, 0.0], [3.0, 0.0, 0.0]])
    out = gather_tof_table(t, 2, 2)
    assert out.shape == (2, 2) and torch.equal(out, t[:, :2])
    z = gather_tof_table(torch.zeros((2, 4)), 0, 2)
    assert z.shape == (2, 1, 1)


# ---- training-step exchange (stofnet_b200/train.py): flat gradient bucket all-reduce + global blur max -------------
def _loss_and_grads(sd, x, gt, blur_max_value, lam=1e-2):
    """Stand-in for the per-rank CUDA compute: the oracle's torch port + torch autograd on CPU (main.py:228-232 with
    the target normalised by a GIVEN batch-wide maximum)."""
    import torch.nn.functional as F
    from oracle import torch_port
    from stofnet_b200.utils.gaussian import gaussian_kernel
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    pred = torch_port.stofnet_forward(params, x)
    mask = torch.zeros_like(pred)
    mask.scatter_(2, gt.clamp(min=0), 1)
    mask[..., :1] = 0
    gk = torch.tensor(gaussian_kernel(7, 1.0), dtype=torch.float32).view(1, 1, -1)
    blur = F.conv1d(mask, gk, padding=3)
    local_max = blur.max()
    mx = local_max if blur_max_value is None else blur_max_value
    blur = blur / mx * 20
    loss = F.mse_loss(pred, blur) + lam * pred.abs().mean()
    loss.backward()
    flat = torch.cat([params[k].grad.reshape(-1) for k in sd])
    return loss.detach(), flat, local_max.detach()


def _train_inputs():
    from stofnet_b200.synth import synth_waveforms
    sd = {k: torch.from_numpy(v) for k, v in load_npz("weights_clean-serenity.npz").items()}
    x = torch.from_numpy(synth_waveforms(4, 160, seed=9, rf=10)).unsqueeze(1)
    # rank 1's rows hold two adjacent echoes (blur max > single tap): the batch-wide maximum lives on one shard only
    gt = torch.tensor([[[100, 0]], [[300, 0]], [[200, 201]], [[50, 400]]], dtype=torch.long)
    return sd, x, gt


def _train_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from stofnet_b200.train import average_gradients, global_blur_max
        sd, x, gt = _train_inputs()
        lo, hi = shard_rows(x.shape[0], world, rank)
        _, _, local_max = _loss_and_grads(sd, x[lo:hi], gt[lo:hi], None)
        mx = global_blur_max(local_max.reshape(1).clone())
        loss, flat, _ = _loss_and_grads(sd, x[lo:hi], gt[lo:hi], mx[0])
        average_gradients(flat)
        dist.all_reduce(loss)
        np.save(os.path.join(out_dir, "grads_%d.npy" % rank), flat.numpy())
        np.save(os.path.join(out_dir, "loss_%d.npy" % rank), (loss / world).numpy())
    finally:
        dist.destroy_process_group()


def test_data_parallel_gradients_match_undivided_batch(tmp_path):
    world = 2
    mp.spawn(_train_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    sd, x, gt = _train_inputs()
    loss, flat, _ = _loss_and_grads(sd, x, gt, None)
    g0 = np.load(os.path.join(str(tmp_path), "grads_0.npy"))
    g1 = np.load(os.path.join(str(tmp_path), "grads_1.npy"))
    assert np.array_equal(g0, g1)                                   # every rank ends with the same bucket
    assert np.abs(g0 - flat.numpy()).max() <= 1e-5 * np.abs(flat.numpy()).max()
    assert abs(float(np.load(os.path.join(str(tmp_path), "loss_0.npy"))) - float(loss)) <= 1e-5 * abs(float(loss))


def test_cosine_lr_matches_torch_scheduler():
    from stofnet_b200.train import cosine_lr
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.AdamW([p], lr=5e-4)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, 80)
    for e in range(10):
        assert abs(opt.param_groups[0]["lr"] - cosine_lr(5e-4, e, 80)) < 1e-12
        opt.step()
        sch.step()
