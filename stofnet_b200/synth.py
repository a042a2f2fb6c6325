"""Deterministic synthetic echo waveforms.

The reference's datasets are absent (stof_chirp101 blob and the PALA submodule are
missing), so every input in tests and in bench.py is synthetic: per row a few
Gaussian-windowed tone bursts plus noise, then the reference's ``NormalizeVol``
(divide by the row abs-max, /root/reference/utils/transforms.py:11-18).

numpy only (PCG64 streams are platform independent).  The float32 generators evaluate exp / sin
in float32, whose last bit depends on the SIMD path numpy picks for the host CPU; ``q15=True``
evaluates in float64 and rounds to the 2^-15 grid of a 16-bit ADC sample, which makes the
waveform bit-identical on any host (a float64 last-bit difference survives the rounding with
probability ~2^-38 per sample).  The at-scale parity fixtures (tests/golden/scale_*.npz) hold
only hashes and outputs and rely on that; every fixture carries the SHA-256 of its inputs.
"""
import numpy as np

__all__ = ["synth_waveforms", "valid_length", "pala_shaped"]


def valid_length(L: int) -> bool:
    """Input-length constraint of the reference forward with the semi-global block:
    L >= 80 and (L mod 80) even (/root/reference/models/stofnet.py:91,111-115)."""
    return L >= 80 and (L % 80) % 2 == 0


def _q15(x):
    """float64 rows in [-1, 1] -> float32 on the 2^-15 grid (what a 16-bit ADC delivers)."""
    return (np.rint(x * 32767.0) * (1.0 / 32768.0)).astype(np.float32)


def synth_waveforms(N: int, L: int, seed: int = 3008, rf: int = 10, max_echoes: int = 5,
                    noise: float = 0.01, chunk: int = 1024, q15: bool = False, return_centres: bool = False):
    """(N, L) float32 rows in [-1, 1]; 1..max_echoes echoes per row.

    echo = a * exp(-0.5 ((t - c) / sigma)^2) * sin(2 pi (t - c) / T),
    c ~ U(0.1 L, 0.9 L), a ~ U(0.2, 1), sigma = 4 rf, T = 2.5 rf samples.
    q15: host-independent variant (float64 evaluation, 2^-15 grid); a different stream of values than q15=False.
    return_centres: also return the (N, max_echoes) float32 echo centres c (0 = absent echo): the ground-truth
    time-of-arrival table in the layout toa_rmse expects (zeros are "no echo").
    """
    ft = np.float64 if q15 else np.float32
    rng = np.random.default_rng(seed)
    out = np.empty((N, L), dtype=np.float32)
    centres = np.zeros((N, max_echoes), dtype=np.float32)
    t = np.arange(L, dtype=ft)[None, :]
    sigma = ft(4.0 * rf)
    period = ft(2.5 * rf)
    for s in range(0, N, chunk):
        n = min(chunk, N - s)
        k = rng.integers(1, max_echoes + 1, size=n)
        c = rng.uniform(0.1 * L, 0.9 * L, size=(n, max_echoes)).astype(ft)
        a = rng.uniform(0.2, 1.0, size=(n, max_echoes)).astype(ft)
        a *= (np.arange(max_echoes)[None, :] < k[:, None])
        x = rng.normal(0.0, noise, size=(n, L)).astype(ft)
        for e in range(max_echoes):
            d = t - c[:, e:e + 1]
            x += a[:, e:e + 1] * np.exp(ft(-0.5) * (d / sigma) ** 2) * np.sin(ft(2 * np.pi) * d / period)
        x /= np.abs(x).max(axis=1, keepdims=True)
        out[s:s + n] = _q15(x) if q15 else x
        centres[s:s + n] = np.where(a > 0, c, 0).astype(np.float32)
    return (out, centres) if return_centres else out


def pala_shaped(channels: int, frames: int, L: int, seed: int = 3008, rf: int = 10,
                noise: float = 0.01, q15: bool = False) -> np.ndarray:
    """(frames*channels, L) float32: per frame 3..10 scatterers whose echo centre follows
    a hyperbolic delay across the channel aperture (PALA plane-wave RF shape, config 3).
    Rows are frame-major then channel, i.e. the reference's fold of the channel dim into
    the batch dim (/root/reference/main.py:300-302)."""
    ft = np.float64 if q15 else np.float32
    rng = np.random.default_rng(seed)
    out = np.empty((frames * channels, L), dtype=np.float32)
    t = np.arange(L, dtype=ft)[None, :]
    ch = (np.arange(channels, dtype=ft) - (channels - 1) / 2)[:, None]
    sigma = ft(4.0 * rf)
    period = ft(2.5 * rf)
    for f in range(frames):
        ns = int(rng.integers(3, 11))
        depth = rng.uniform(0.15 * L, 0.8 * L, size=ns).astype(ft)
        lat = rng.uniform(-0.4 * channels, 0.4 * channels, size=ns).astype(ft)
        amp = rng.uniform(0.2, 1.0, size=ns).astype(ft)
        x = rng.normal(0.0, noise, size=(channels, L)).astype(ft)
        pitch = ft(0.004 * L)  # samples of delay per channel pitch of lateral offset
        for s in range(ns):
            c = 0.5 * (depth[s] + np.sqrt(depth[s] ** 2 + ((ch - lat[s]) * pitch) ** 2))
            d = t - c
            x += amp[s] * np.exp(ft(-0.5) * (d / sigma) ** 2) * np.sin(ft(2 * np.pi) * d / period)
        x /= np.abs(x).max(axis=1, keepdims=True)
        out[f * channels:(f + 1) * channels] = _q15(x) if q15 else x
    return out
