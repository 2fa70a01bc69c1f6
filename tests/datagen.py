"""Seeded synthetic inputs shared by the parity tests, the golden-vector generator and bench.py
(binary / float descriptor sets with planted neighbours and adversarial rows)."""
import numpy as np


def binary_descriptors(nq, nt, nbytes, seed, adversarial=True):
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, (nt, nbytes), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, nbytes), dtype=np.uint8)
    half = nq // 2
    if half and nt:
        src = rng.integers(0, nt, half)
        flips = rng.random((half, nbytes * 8)) < 0.08
        q[:half] = t[src] ^ np.packbits(flips, axis=1)
    if adversarial and nt >= 8 and nq >= 8:
        t[3] = t[1]                       # exact duplicate train rows -> (dist, idx) tie, lowest index wins
        q[-1] = t[1]                      # distance 0 to both
        q[-2] = t[5]; q[-2, -1] ^= 0xFF   # differs from a train row only in the last byte (BRUTEFORCENMS ignores it)
        q[-3] = t[6]; q[-3, -2] ^= 0x0F   # differs only in the second-to-last byte
        q[-4] = 0; q[-5] = 255
    return q, t


def sift_like(nq, nt, dim, seed, integer=True):
    rng = np.random.default_rng(seed)

    def gen(n):
        x = np.abs(rng.standard_normal((n, dim)))
        x /= np.linalg.norm(x, axis=1, keepdims=True) + 1e-12
        if integer:
            return np.minimum(np.round(x * 512.0), 255.0).astype(np.float32)
        return x.astype(np.float32)
    t = gen(nt)
    q = gen(nq)
    half = nq // 2
    if half and nt:
        src = rng.integers(0, nt, half)
        noise = rng.standard_normal((half, dim)) * (0.05 * (512.0 / np.sqrt(dim) if integer else 1.0 / np.sqrt(dim)))
        p = t[src] + noise
        q[:half] = (np.clip(np.round(p), 0, 255) if integer else p).astype(np.float32)
    if nt >= 4 and nq >= 2:
        t[2] = t[0]                       # duplicate rows -> tie
        q[-1] = t[0]
    return q, t


def signed_float(nq, nt, dim, seed):
    """SURF/KAZE-like: unit-norm signed float32 (non-integer values)."""
    rng = np.random.default_rng(seed)
    t = rng.standard_normal((nt, dim)); t /= np.linalg.norm(t, axis=1, keepdims=True)
    q = rng.standard_normal((nq, dim)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    half = nq // 2
    if half and nt:
        src = rng.integers(0, nt, half)
        p = t[src] + rng.standard_normal((half, dim)) * 0.05 / np.sqrt(dim)
        q[:half] = p / np.linalg.norm(p, axis=1, keepdims=True)
    return q.astype(np.float32), t.astype(np.float32)
