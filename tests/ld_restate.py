"""numpy restatement of the device `ldsampler` (csrc/shading.cuh ld_next1D / ld_next2D) — test infrastructure.
radical_inverse2 / sobol2 follow include/mitsuba/core/qmc.h:43-59,82-87 and are pinned bit for bit on tests/golden/qmc.npz (made by
the reference); the Philox keying and the hashed permutation are the device's own and are pinned against its dump."""
import numpy as np

U32 = np.uint32
M32 = np.uint64(0xffffffff)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c = [np.asarray(x, np.uint64) & M32 for x in np.broadcast_arrays(c0, c1, c2, c3)]
    k0, k1 = np.uint64(k0), np.uint64(k1)
    for _ in range(10):
        p0, p1 = np.uint64(0xD2511F53) * c[0], np.uint64(0xCD9E8D57) * c[2]
        c = [((p1 >> np.uint64(32)) ^ c[1] ^ k0) & M32, p1 & M32, ((p0 >> np.uint64(32)) ^ c[3] ^ k1) & M32, p0 & M32]
        k0, k1 = (k0 + np.uint64(0x9E3779B9)) & M32, (k1 + np.uint64(0xBB67AE85)) & M32
    return [x.astype(U32) for x in c]


def radical_inverse2(n, scramble):
    n = np.asarray(n, np.uint64) & M32
    rev = np.zeros_like(n)
    for b in range(32):
        rev |= ((n >> np.uint64(b)) & np.uint64(1)) << np.uint64(31 - b)
    v = (rev >> np.uint64(8)) ^ (np.asarray(scramble, np.uint64) & np.uint64(0xffffff))
    return (v.astype(np.float32) / np.float32(16777216.0)).astype(np.float32)


def sobol2_bits(n, scramble):
    n = np.asarray(n, np.uint64) & M32
    s = np.asarray(scramble, np.uint64) & M32
    s = np.broadcast_to(s, n.shape).copy()
    v = np.uint64(1 << 31)
    for _ in range(32):
        s = np.where(n & np.uint64(1), s ^ v, s)
        n = n >> np.uint64(1)
        v = v ^ (v >> np.uint64(1))
    return s


def sobol2(n, scramble, clamp=False):
    """(float) scramble / (float) (1ULL << 32); the device clamps the result below 1 (the division rounds up for the top 128 values)"""
    out = (sobol2_bits(n, scramble).astype(np.float32) / np.float32(4294967296.0)).astype(np.float32)
    return np.minimum(out, np.float32(0.99999994)) if clamp else out


def permute(i, w, p):
    i = np.asarray(i, np.uint64) & M32
    w, p = np.uint64(w), np.asarray(p, np.uint64) & M32
    mul = lambda a, b: (a * np.uint64(b)) & M32
    i = i ^ p; i = mul(i, 0xe170893d); i ^= p >> np.uint64(16); i ^= (i & w) >> np.uint64(4); i ^= p >> np.uint64(8); i = mul(i, 0x0929eb3f)
    i ^= p >> np.uint64(23); i ^= (i & w) >> np.uint64(1); i = (i * (np.uint64(1) | (p >> np.uint64(27)))) & M32; i = mul(i, 0x6935fa69)
    i ^= (i & w) >> np.uint64(11); i = mul(i, 0x74dcb303); i ^= (i & w) >> np.uint64(2); i = mul(i, 0x9e501cc3); i ^= (i & w) >> np.uint64(2)
    i = mul(i, 0xc860a3df); i &= w; i ^= i >> np.uint64(5)
    return ((i + p) & w).astype(U32)


def u01(word):
    return ((np.asarray(word, U32) >> U32(9)) | U32(0x3f800000)).view(np.float32) - np.float32(1.0)


def dump(px, py, first, n, draws, seed, dimension, log2n):
    """What mtsb_render_sampler_dump returns for the ldsampler: [n, draws, 3]"""
    s = (first + np.arange(n)).astype(np.uint64)
    mask = (1 << log2n) - 1
    out = np.zeros((n, draws, 3), np.float32)
    indep = 0          # count of independent draws so far (same for every sample)

    def independent():
        nonlocal indep
        c = indep
        indep += 1
        w = philox4x32_10(px, py, s, 0x80000000 | (c >> 2), seed, 0x6d747362)
        return u01(w[c & 3])
    for k in range(draws):
        if k < dimension:
            w = philox4x32_10(px, py, 16 + k, 0x6c643244, seed, 0x6d747362)
            i = permute(s & np.uint64(mask), mask, w[2])
            out[:, k, 0] = radical_inverse2(i, w[0]); out[:, k, 1] = sobol2(i, w[1], clamp=True)
        else:
            out[:, k, 0] = independent(); out[:, k, 1] = independent()
    for k in range(draws):
        if k < dimension:
            w = philox4x32_10(px, py, k, 0x6c643144, seed, 0x6d747362)
            out[:, k, 2] = radical_inverse2(permute(s & np.uint64(mask), mask, w[2]), w[0])
        else:
            out[:, k, 2] = independent()
    return out
