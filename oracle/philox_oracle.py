"""numpy restatement of csrc/philox.cuh (Philox4x32-10, Salmon et al. SC'11).  TEST INFRASTRUCTURE ONLY.

The reference draws batch indices with numpy's global Mersenne Twister inside Dataset.sample (fql submodule;
call site /root/reference/task/offline_task_simulated.py:53-58) and noise with jax.random (threefry) inside
FQLAgent.update / sample_actions. Neither stream can be reproduced on the device, so the product uses one
counter-based generator for every draw; this file pins that generator bit for bit: integer words and row
indices must match exactly, uniforms exactly (they are exact fp32 functions of the words), normals to fp32
round-off of logf/cospif.

Known-answer test: the Random123 reference vector philox4x32-10(counter=0, key=0) =
6627e8d5 e169c58d bc57ac4c 9b00dbd8 (checked in tests/test_oracle_cpu.py).
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(seed, step, row, slot):
    """Vectorised over numpy-broadcastable integer arrays. key = (seed lo, seed hi), counter = (step lo, step hi,
    row, slot). Returns four uint32 arrays."""
    seed, step, row, slot = np.broadcast_arrays(np.asarray(seed, np.uint64), np.asarray(step, np.uint64),
                                                np.asarray(row, np.uint64), np.asarray(slot, np.uint64))
    k0, k1 = seed & MASK, seed >> np.uint64(32)
    c0, c1, c2, c3 = step & MASK, step >> np.uint64(32), row & MASK, slot & MASK
    for r in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0 = (k0 + np.uint64(W0)) & MASK
        k1 = (k1 + np.uint64(W1)) & MASK
    return tuple(x.astype(np.uint32) for x in (c0, c1, c2, c3))


def uniform(seed, step, row, slot):
    x, _, _, _ = philox4x32_10(seed, step, row, slot)
    return ((x >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)


def index(seed, step, row, slot, n):
    x, _, _, _ = philox4x32_10(seed, step, row, slot)
    return ((x.astype(np.uint64) * np.uint64(n)) >> np.uint64(32)).astype(np.int64)


def normal(seed, step, row, slot):
    x, y, _, _ = philox4x32_10(seed, step, row, slot)
    u1 = (((x >> np.uint32(8)) + np.uint32(1)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)
    u2 = ((y >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)
    return (np.sqrt(np.float32(-2.0) * np.log(u1)) * np.cos(np.float32(2.0 * np.pi) * u2.astype(np.float64))).astype(
        np.float32)


def batch_indices(seed, step, batch_size, action_dim, n_rows):
    """Row indices the prep kernel gathers for one member at lockstep step `step` (slot 4A+1)."""
    return index(seed, step, np.arange(batch_size), 4 * action_dim + 1, n_rows)
