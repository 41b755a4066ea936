"""Seeded synthetic 44.1 kHz mono test audio.

`make_track` (numpy, host) is the generator shared by the parity tests, the golden vectors and the CPU
baseline; `make_batch_torch` draws a statistically equivalent batch directly on the GPU for the
throughput benchmark (1024 x 30 s tracks are far too slow to synthesise on the host).

Track recipe: 8 random sinusoids (50-8000 Hz, amplitude U(500, 4000)) + N(0, 300^2) noise + 20 random
exponentially decaying noise bursts (so that peak picking has real onsets), clipped to int16.
"""
import numpy as np

SAMPLE_RATE = 44100


def make_track(seed, seconds=30.0, sample_rate=SAMPLE_RATE, dtype=np.int16, num_samples=None):
    rng = np.random.default_rng(seed)
    n = int(round(seconds * sample_rate)) if num_samples is None else int(num_samples)
    t = np.arange(n, dtype=np.float64) / sample_rate
    x = np.zeros(n, dtype=np.float64)
    freqs = rng.uniform(50., 8000., 8)
    amps = rng.uniform(500., 4000., 8)
    phases = rng.uniform(0., 2 * np.pi, 8)
    for f, a, p in zip(freqs, amps, phases):
        x += a * np.sin(2 * np.pi * f * t + p)
    x += rng.normal(0., 300., n)
    n_bursts = 20
    starts = rng.integers(0, max(n - 1, 1), n_bursts)
    burst_amp = rng.uniform(2000., 9000., n_bursts)
    decay = rng.uniform(0.01, 0.08, n_bursts) * sample_rate
    for s, a, d in zip(starts, burst_amp, decay):
        ln = int(min(n - s, 6 * d))
        if ln <= 0:
            continue
        env = a * np.exp(-np.arange(ln) / d)
        x[s:s + ln] += env * rng.normal(0., 1., ln)
    x = np.clip(np.round(x), -32768, 32767).astype(np.int16)
    if np.dtype(dtype) == np.int16:
        return x
    # float variant: same samples rescaled like madmom's Signal(dtype=float) (signal.py:266-292)
    return (x.astype(np.float64) / 32767.).astype(dtype)


def make_batch_torch(n_tracks, seconds=30.0, sample_rate=SAMPLE_RATE, device="cuda", seed=1000,
                     dtype=None):
    """(n_tracks, num_samples) int16 tensor generated on `device` with the recipe above."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    n = int(round(seconds * sample_rate))
    out = torch.empty((n_tracks, n), dtype=torch.int16, device=device)
    t = torch.arange(n, device=device, dtype=torch.float32) / sample_rate
    blk = 64
    for b0 in range(0, n_tracks, blk):
        nb = min(blk, n_tracks - b0)
        x = torch.randn((nb, n), device=device, generator=g) * 300.
        freqs = 50. + torch.rand((nb, 8), device=device, generator=g) * 7950.
        amps = 500. + torch.rand((nb, 8), device=device, generator=g) * 3500.
        ph = torch.rand((nb, 8), device=device, generator=g) * (2 * np.pi)
        for k in range(8):
            x += amps[:, k:k + 1] * torch.sin(2 * np.pi * freqs[:, k:k + 1] * t[None, :] + ph[:, k:k + 1])
        # bursts: 20 per track, decaying noise
        starts = (torch.rand((nb, 20), device=device, generator=g) * (n - 1)).long()
        bamp = 2000. + torch.rand((nb, 20), device=device, generator=g) * 7000.
        decay = (0.01 + torch.rand((nb, 20), device=device, generator=g) * 0.07) * sample_rate
        ln = int(0.5 * sample_rate)
        ar = torch.arange(ln, device=device, dtype=torch.float32)
        for k in range(20):
            env = bamp[:, k:k + 1] * torch.exp(-ar[None, :] / decay[:, k:k + 1])
            noise = torch.randn((nb, ln), device=device, generator=g)
            idx = starts[:, k:k + 1] + torch.arange(ln, device=device)[None, :]
            # a burst that runs past the end of the track is cut off: its tail is zeroed before the indices are
            # clamped, so the colliding atomic adds on the last sample all add 0.0 and the result does not depend on
            # their order (the same seed then gives the same samples on every rank and in every run)
            burst = (env * noise) * (idx < n)
            x.scatter_add_(1, idx.clamp_(max=n - 1), burst)
        out[b0:b0 + nb] = x.round_().clamp_(-32768, 32767).to(torch.int16)
    if dtype is not None and dtype != torch.int16:
        return (out.to(torch.float32) / 32767.).to(dtype)
    return out
