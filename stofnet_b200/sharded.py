"""Row-sharded inference across the GPUs of one box.

The reference is single-GPU (`#SBATCH --gres=gpu:1` in every script); its only "parallelism" is folding
PALA channels into the batch dimension (main.py:209,301).  Rows (waveforms) are fully independent -
Conv1d / max-pool / NMS act along the sample axis only - so N GPUs each take a contiguous block of rows,
weights are replicated, and there is NO collective on the data path.  The one exchange is the final
time-of-flight gather, because the shape of the reference's table is a property of the WHOLE batch:

    coords = zeros((N, K_max)),  K_max = max_k counts_k          utils/mask2samples.py:92-99
    zeros((N, 1, 1)) when nothing is detected anywhere           utils/mask2samples.py:87-88

-> one all_reduce(MAX) of the per-rank K_max, one all_gather of the fixed-stride (rows, K_max) table.

One process per GPU (torchrun); `torch.distributed` with the NCCL backend over NVLink on the GPU box, gloo
in the CPU tests of the host logic (tests/test_sharded.py).
"""
import torch
import torch.distributed as dist

__all__ = ["shard_rows", "gather_tof_table", "infer_sharded", "TofGatherPipeline"]


def shard_rows(n_rows, world_size, rank):
    """[start, stop) of this rank's contiguous row block; the first n_rows % world_size ranks get one
    extra row, so any N works and concatenating the blocks in rank order restores the batch order."""
    if n_rows < 0 or world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("shard_rows: n_rows=%d world_size=%d rank=%d" % (n_rows, world_size, rank))
    base, extra = divmod(n_rows, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def gather_tof_table(local_coords, local_kmax, n_rows_total, group=None, channels=1):
    """Assemble the reference-shaped global ToF table from per-rank tables.

    local_coords  (n_local, cap) float32, zero padded right, cap >= local_kmax (this rank's rows in order;
                  what stof_peaks / mask2coords produced locally, already divided by upsample_factor)
    local_kmax    max detections in any local row
    returns       (n_rows_total, K_max) on every rank, or zeros((n_rows_total, channels, 1)) when no rank
                  detected anything - exactly what mask2coords returns on the undivided batch.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = local_coords.device
    lo, hi = shard_rows(n_rows_total, world, rank)
    if local_coords.shape[0] != hi - lo:
        raise ValueError("gather_tof_table: rank %d holds %d rows, its shard of %d is [%d, %d)"
                         % (rank, local_coords.shape[0], n_rows_total, lo, hi))
    kmax = torch.tensor([int(local_kmax)], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(kmax, op=dist.ReduceOp.MAX, group=group)
    k = int(kmax.item())                     # the one host sync (the reference has it too: int(max(counts)), :93)
    if k == 0:
        return torch.zeros((n_rows_total, channels, 1), dtype=torch.float32, device=dev)
    # fixed stride K_max: narrow or zero-extend the local table
    if local_coords.shape[1] >= k:
        send = local_coords[:, :k].contiguous()
    else:
        send = torch.zeros((local_coords.shape[0], k), dtype=torch.float32, device=dev)
        send[:, :local_coords.shape[1]] = local_coords
    if world == 1:
        return send
    out = torch.empty((n_rows_total, k), dtype=torch.float32, device=dev)
    if n_rows_total % world == 0:
        dist.all_gather_into_tensor(out, send, group=group)          # equal shards: one flat collective
    else:
        # ragged shards: collectives need equal sizes, so every rank sends ceil(N / world) rows (the short
        # shards zero-extend by one row) and the table is compacted from the padded gather
        rows_max = -(-n_rows_total // world)
        padded = torch.zeros((rows_max, k), dtype=torch.float32, device=dev)
        padded[:send.shape[0]] = send
        stage = torch.empty((world * rows_max, k), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(stage, padded, group=group)
        for r in range(world):
            a, b = shard_rows(n_rows_total, world, r)
            out[a:b] = stage[r * rows_max:r * rows_max + (b - a)]
    return out


def infer_sharded(model, frame_local, n_rows_total, window_size, threshold=None, upsample_factor=None,
                  group=None, cap=64):
    """Eval step of main.py:314,320 on this rank's rows + the global ToF gather.

    frame_local (n_local, 1, L) CUDA tensor = rows shard_rows(n_rows_total, world, rank) of the batch.
    Returns (masks_pred_local, coords_global)."""
    from .utils.mask2samples import _as_rows, _run_peaks
    r = model.upsample_factor if upsample_factor is None else upsample_factor
    with torch.no_grad():
        masks = model(frame_local)
    rows = _as_rows(masks)
    int_r = isinstance(r, int) or float(r).is_integer()
    coords, _, _, kmax, _ = _run_peaks(rows, window_size, threshold, int(r) if int_r else 1, False, cap)
    if not int_r:
        coords = coords / r
    table = gather_tof_table(coords, kmax, n_rows_total, group=group,
                             channels=masks.shape[1] if masks.dim() == 3 else 1)
    return masks, table


class TofGatherPipeline:
    """The per-step ToF gather taken off the critical path (streaming inference over many batches).

    gather_tof_table() is exact but serial: MAX all-reduce -> host reads K_max -> all-gather, all on the compute stream, and
    the all-gather moves the full cap-wide table.  In a stream of batches the table of step i is only needed once step i+1
    is already running, so here
      * `submit` copies the first `k_opt` columns of the local table into a slot buffer (k_opt = the previous steps' global
        K_max, the optimistic stride; 1/2 .. 1/4 of `cap` in practice) and issues the MAX all-reduce + all-gather on a SIDE
        stream behind an event, so the next batch's forward overlaps the exchange;
      * `collect` (called one step later, or at the end) waits for the slot, reads the global K_max, and returns the
        reference-shaped table; when a batch turns out to need more columns than k_opt the caller's full local table of
        that step (kept alive by double buffering: `submit` returns the slot index) is re-gathered at the exact width - the
        fallback costs one serial gather and raises k_opt.
    Equal shards only (rows_local the same on every rank: what the bench and the 128 x 1000 PALA fold have).
    CPU tensors / gloo: the same logic runs inline without streams (tests/test_sharded.py)."""

    def __init__(self, rows_local, cap, device, group=None, k_init=None, slots=2):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rows, self.cap, self.device = int(rows_local), int(cap), device
        self.k_seen = int(k_init) if k_init else 0        # largest global K_max seen so far
        self.k_opt = self._stride(self.k_seen) if k_init else self.cap     # first step: full width
        self.cuda = device.type == "cuda"
        self.stream = torch.cuda.Stream(device) if self.cuda else None
        self.slots = slots
        self.send = [torch.zeros((self.rows * self.cap,), dtype=torch.float32, device=device) for _ in range(slots)]
        self.recv = [torch.zeros((self.world * self.rows * self.cap,), dtype=torch.float32, device=device) for _ in range(slots)]
        self.kmax = [torch.zeros((1,), dtype=torch.int64, device=device) for _ in range(slots)]
        self.done = [torch.cuda.Event() if self.cuda else None for _ in range(slots)]
        self.k_used = [0] * slots
        self.i = 0
        self.fallbacks = 0

    def _stride(self, k_seen):
        """optimistic stride for the next steps: the largest K_max seen plus 1/8 headroom (at least 4 columns)"""
        return min(self.cap, max(1, k_seen + max(4, k_seen // 8)))

    def submit(self, local_coords, local_kmax):
        """local_coords (rows, cap) fp32 as written by stof_peaks; local_kmax: 1-element int64 tensor on the device (stats[:1])
        or an int.  Returns the slot; the caller must keep `local_coords` of this step intact until collect(slot)."""
        slot = self.i % self.slots
        self.i += 1
        k = min(self.k_opt, self.cap)
        self.k_used[slot] = k
        send = self.send[slot][:self.rows * k].view(self.rows, k)
        send.copy_(local_coords[:, :k])                                   # tiny; on the compute stream
        if isinstance(local_kmax, torch.Tensor):
            self.kmax[slot].copy_(local_kmax.reshape(-1)[:1])
        else:
            self.kmax[slot].fill_(int(local_kmax))
        recv = self.recv[slot][:self.world * self.rows * k].view(self.world * self.rows, k)
        if self.cuda:
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ready)
                self._exchange(self.kmax[slot], send, recv)
                self.done[slot].record(self.stream)
        else:
            self._exchange(self.kmax[slot], send, recv)
        return slot

    def _exchange(self, kmax, send, recv):
        if self.world > 1:
            dist.all_reduce(kmax, op=dist.ReduceOp.MAX, group=self.group)
            dist.all_gather_into_tensor(recv, send, group=self.group)
        else:
            recv.copy_(send)

    def collect(self, slot, local_coords=None, channels=1):
        """Global (world * rows, K_max) table of the step submitted into `slot` (zeros((N, channels, 1)) when no rank
        detected anything, utils/mask2samples.py:87-88).  local_coords: that step's full local table, needed only when
        the optimistic stride was too narrow."""
        if self.cuda:
            self.done[slot].synchronize()
        k, kg = self.k_used[slot], int(self.kmax[slot].item())
        n_total = self.world * self.rows
        if kg == 0:
            return torch.zeros((n_total, channels, 1), dtype=torch.float32, device=self.device)
        self.k_seen = max(self.k_seen, kg)
        self.k_opt = self._stride(self.k_seen)
        if kg > k:
            if local_coords is None:
                raise RuntimeError("TofGatherPipeline: K_max=%d exceeds the optimistic stride %d and the local table was not kept" % (kg, k))
            self.fallbacks += 1
            return gather_tof_table(local_coords, kg, n_total, group=self.group, channels=channels)
        return self.recv[slot][:n_total * k].view(n_total, k)[:, :kg]
