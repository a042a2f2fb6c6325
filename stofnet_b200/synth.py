"""Deterministic synthetic echo waveforms.

The reference's datasets are absent (stof_chirp101 blob and the PALA submodule are
missing), so every input in tests and in bench.py is synthetic: per row a few
Gaussian-windowed tone bursts plus noise, then the reference's ``NormalizeVol``
(divide by the row abs-max, /root/reference/utils/transforms.py:11-18).

numpy only (PCG64 streams are platform independent), so the same seed gives the same
waveforms in the build container and on the GPU box.
"""
import numpy as np

__all__ = ["synth_waveforms", "valid_length", "pala_shaped"]


def valid_length(L: int) -> bool:
    """Input-length constraint of the reference forward with the semi-global block:
    L >= 80 and (L mod 80) even (/root/reference/models/stofnet.py:91,111-115)."""
    return L >= 80 and (L % 80) % 2 == 0


def synth_waveforms(N: int, L: int, seed: int = 3008, rf: int = 10, max_echoes: int = 5,
                    noise: float = 0.01, chunk: int = 1024) -> np.ndarray:
    """(N, L) float32 rows in [-1, 1]; 1..max_echoes echoes per row.

    echo = a * exp(-0.5 ((t - c) / sigma)^2) * sin(2 pi (t - c) / T),
    c ~ U(0.1 L, 0.9 L), a ~ U(0.2, 1), sigma = 4 rf, T = 2.5 rf samples.
    """
    rng = np.random.default_rng(seed)
    out = np.empty((N, L), dtype=np.float32)
    t = np.arange(L, dtype=np.float32)[None, :]
    sigma = np.float32(4.0 * rf)
    period = np.float32(2.5 * rf)
    for s in range(0, N, chunk):
        n = min(chunk, N - s)
        k = rng.integers(1, max_echoes + 1, size=n)
        c = rng.uniform(0.1 * L, 0.9 * L, size=(n, max_echoes)).astype(np.float32)
        a = rng.uniform(0.2, 1.0, size=(n, max_echoes)).astype(np.float32)
        a *= (np.arange(max_echoes)[None, :] < k[:, None])
        x = rng.normal(0.0, noise, size=(n, L)).astype(np.float32)
        for e in range(max_echoes):
            d = t - c[:, e:e + 1]
            x += a[:, e:e + 1] * np.exp(np.float32(-0.5) * (d / sigma) ** 2) * np.sin(
                np.float32(2 * np.pi) * d / period)
        x /= np.abs(x).max(axis=1, keepdims=True)
        out[s:s + n] = x
    return out


def pala_shaped(channels: int, frames: int, L: int, seed: int = 3008, rf: int = 10,
                noise: float = 0.01) -> np.ndarray:
    """(frames*channels, L) float32: per frame 3..10 scatterers whose echo centre follows
    a hyperbolic delay across the channel aperture (PALA plane-wave RF shape, config 3).
    Rows are frame-major then channel, i.e. the reference's fold of the channel dim into
    the batch dim (/root/reference/main.py:300-302)."""
    rng = np.random.default_rng(seed)
    out = np.empty((frames * channels, L), dtype=np.float32)
    t = np.arange(L, dtype=np.float32)[None, :]
    ch = (np.arange(channels, dtype=np.float32) - (channels - 1) / 2)[:, None]
    sigma = np.float32(4.0 * rf)
    period = np.float32(2.5 * rf)
    for f in range(frames):
        ns = int(rng.integers(3, 11))
        depth = rng.uniform(0.15 * L, 0.8 * L, size=ns).astype(np.float32)
        lat = rng.uniform(-0.4 * channels, 0.4 * channels, size=ns).astype(np.float32)
        amp = rng.uniform(0.2, 1.0, size=ns).astype(np.float32)
        x = rng.normal(0.0, noise, size=(channels, L)).astype(np.float32)
        pitch = np.float32(0.004 * L)  # samples of delay per channel pitch of lateral offset
        for s in range(ns):
            c = 0.5 * (depth[s] + np.sqrt(depth[s] ** 2 + ((ch - lat[s]) * pitch) ** 2))
            d = t - c
            x += amp[s] * np.exp(np.float32(-0.5) * (d / sigma) ** 2) * np.sin(
                np.float32(2 * np.pi) * d / period)
        x /= np.abs(x).max(axis=1, keepdims=True)
        out[f * channels:(f + 1) * channels] = x
    return out
