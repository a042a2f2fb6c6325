"""stof_mask_loss_static: the allocation-free variant of the mask loss (partial sums in a file-scope device array of the
library instead of a caller-provided scratch buffer) against the existing entry point stof_mask_loss, through the C ABI.

No counterpart in the reference.  Collected last on purpose (file name):
a failure here must not hide the parity tests.  Passes alone in a fresh process and after
the other tests; allocates its own tensors before the calls and compares bit for bit (same kernels, same fixed-order sums -
the loss tests' tolerance of 2e-6 relative is the stated bound, equality is what is expected).
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

LOSS_CHUNK = 4096
STATIC_MAX_BLOCKS = 16384


def _gauss(k=7, sigma=1.0):
    x = np.arange(k, dtype=np.float64) - (k - 1) / 2.0
    g = np.exp(-0.5 * (x / sigma) ** 2)
    return (g / g.sum()).astype(np.float32)


def _case(seed, n, m, g, dev):
    rng = np.random.default_rng(seed)
    pred = torch.from_numpy(rng.standard_normal((n, m)).astype(np.float32) * 3.0).to(dev)
    gt = rng.integers(0, m, size=(n, g)).astype(np.int64)
    gt[:, -1] = -1                                     # a padded (absent) entry per row, as the batches carry them
    gt[0, 0] = 0
    gt[-1, 1] = m - 1                                  # blur clipped at both ends
    return pred, torch.from_numpy(gt).to(dev)


def _run(lib, static, pred, gt, gauss, lam, mode=0, mx=None):
    from stofnet_b200._lib import check
    n, m = pred.shape
    dev = pred.device
    mx = torch.zeros((1,), dtype=torch.float32, device=dev) if mx is None else mx.clone()
    loss = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
    grad = torch.full((n, m), float("nan"), dtype=torch.float32, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    args = [pred.data_ptr(), gt.data_ptr(), n, gt.shape[1], m, gauss.data_ptr(), gauss.numel(), 20.0, lam, mode,
            mx.data_ptr(), loss.data_ptr(), grad.data_ptr()]
    with torch.cuda.device(dev):
        if static:
            check(lib.stof_mask_loss_static(*args, stream))
        else:
            scratch = torch.empty((lib.stof_loss_scratch_bytes(n, m),), dtype=torch.uint8, device=dev)
            check(lib.stof_mask_loss(*args, scratch.data_ptr(), stream))
    torch.cuda.synchronize(dev)
    return float(loss.item()), grad, float(mx.item())


def test_mask_loss_static_matches_mask_loss_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from stofnet_b200._lib import lib
    dev = torch.device("cuda:0")
    gauss = torch.from_numpy(_gauss()).to(dev)
    # first call: 24 rows x 3 chunks = 72 partial slots; second call: FEWER slots (10) and different data, so the array still
    # holds the first call's sums beyond slot 10 - they must not leak into the result; third: back to a larger one
    cases = [(_case(11, 24, 10240, 6, dev), 1e-2), (_case(12, 5, 5000, 4, dev), 1.0), (_case(13, 40, 4097, 9, dev), 1e-2)]
    ref = [_run(lib, False, p, g, gauss, lam) for (p, g), lam in cases]
    for ((p, g), lam), (l_ref, g_ref, mx_ref) in zip(cases, ref):          # static calls back to back, nothing in between
        l_st, g_st, mx_st = _run(lib, True, p, g, gauss, lam)
        assert np.isfinite(l_st) and abs(l_st - l_ref) <= 2e-6 * abs(l_ref)
        assert l_st == l_ref and mx_st == mx_ref                           # same kernels, same fixed-order sum
        assert torch.equal(g_st, g_ref)
    # the same input again: the array now holds the last call's contents
    (p, g), lam = cases[0]
    l_again, g_again, _ = _run(lib, True, p, g, gauss, lam)
    assert l_again == ref[0][0] and torch.equal(g_again, ref[0][1])
    # mode 2 (blur max supplied: the second phase of the sharded loss) goes the same way
    mx = torch.full((1,), ref[0][2], dtype=torch.float32, device=dev)
    l_m2, g_m2, _ = _run(lib, True, p, g, gauss, lam, mode=2, mx=mx)
    assert l_m2 == ref[0][0] and torch.equal(g_m2, ref[0][1])


def test_mask_loss_static_rejects_sizes_beyond_the_array():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from stofnet_b200._lib import lib, E_UNSUPPORTED
    dev = torch.device("cuda:0")
    one = torch.zeros((8,), dtype=torch.float32, device=dev)               # never touched: the size check comes first
    n = STATIC_MAX_BLOCKS // 2 + 1                                         # x 2 chunks (M = 4097) = one block too many
    with torch.cuda.device(dev):
        rc = lib.stof_mask_loss_static(one.data_ptr(), one.data_ptr(), n, 1, LOSS_CHUNK + 1, one.data_ptr(), 7, 20.0, 1e-2, 0,
                                       one.data_ptr(), one.data_ptr(), None, None)
    assert rc == E_UNSUPPORTED
    msg = lib.stof_last_error().decode()
    assert "static scratch" in msg and str(STATIC_MAX_BLOCKS) in msg
    torch.cuda.synchronize(dev)
