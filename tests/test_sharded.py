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
