"""numpy Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11) -- the
counter-based generator the performance-mode sampler runs in-kernel (diner_b200/csrc/common.cuh).  TEST INFRASTRUCTURE:
lets tests hand the parity-mode sampler the very uniforms the performance mode draws.  Pinned by the Random123
known-answer vectors in tests/test_host.py."""
import numpy as np


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """numpy Philox4x32-10 (Salmon et al. 2011), the counter-based generator of the performance-mode sampler."""
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    c0, c1, c2, c3 = (np.asarray(c, np.uint64) for c in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0), np.uint64(k1)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask, lo1, (hi0 ^ c3 ^ k1) & mask, lo0
        k0, k1 = (k0 + np.uint64(0x9E3779B9)) & mask, (k1 + np.uint64(0xBB67AE85)) & mask
    return c0, c1, c2, c3
