"""Counter-based RNG used for every synthetic input:
u(seed, i) = (splitmix64(seed XOR (i * 0x9E3779B97F4A7C15)) >> 11) * 2^-53.
Bit-identical in numpy (host, parity sets) and torch (device, timed sets)."""
import numpy as np

_GOLD = 0x9E3779B97F4A7C15
_M1 = 0xBF58476D1CE4E5B9
_M2 = 0x94D049BB133111EB


def uniform_np(seed, start, n):
    with np.errstate(over="ignore"):
        i = np.arange(start, start + n, dtype=np.uint64)
        z = np.uint64(seed) ^ (i * np.uint64(_GOLD))
        z = z + np.uint64(_GOLD)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(_M1)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(_M2)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)


def _i64(v):
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= (1 << 63) else v


def uniform_torch(seed, start, n, device):
    """Same bits as uniform_np, generated on `device` (int64 wrap-around arithmetic)."""
    import torch

    i = torch.arange(start, start + n, dtype=torch.int64, device=device)
    z = torch.bitwise_xor(torch.full_like(i, _i64(seed)), i * _i64(_GOLD))
    z = z + _i64(_GOLD)

    def lsr(x, k):  # logical shift right on int64
        return torch.bitwise_and(x >> k, (1 << (64 - k)) - 1)

    z = torch.bitwise_xor(z, lsr(z, 30)) * _i64(_M1)
    z = torch.bitwise_xor(z, lsr(z, 27)) * _i64(_M2)
    z = torch.bitwise_xor(z, lsr(z, 31))
    return lsr(z, 11).to(torch.float64) * (2.0 ** -53)
