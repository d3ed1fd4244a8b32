"""Counter-based RNG of the generate mode (oracle; test infrastructure only).

The reference draws noise and timesteps from a torch CPU generator (train_sana.py:183-191,
common/trainer.py:325).  A counter-based stream cannot reproduce mt19937, so bit-parity with
the reference is tested in *inject* mode (noise/u supplied); this module restates the
device's own Philox4x32-10 + Box-Muller stream (yat_b200/csrc/philox.cuh) so the *generate*
mode has a CPU oracle too: the integer stream is exact, timestep normals agree to a few fp32 ulp and
latent normals to ~1e-6 absolute (special-function-unit evaluation on the device).

Layout (must match philox.cuh):
  key     = (seed & 0xffffffff, seed >> 32)
  counter = (quad, sample, offset & 0xffffffff, ((offset >> 32) & 0x7fffffff) | stream << 31)
  stream 0 = latent noise (c0 = oct = element_index // 8 inside the sample; one 32-bit word per
             Box-Muller pair -> 8 normals per call),
  stream 1 = timestep draw (c0 = 0, first normal of the call; two words per pair).
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10; all inputs uint32 arrays (broadcastable). Returns 4 uint32 arrays."""
    c0, c1, c2, c3 = [np.asarray(c, dtype=np.uint64) & MASK for c in (c0, c1, c2, c3)]
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)) & MASK, lo1, (hi0 ^ c3 ^ np.uint64(k1)) & MASK, lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def _counter_hi(offset, stream):
    return ((int(offset) >> 32) & 0x7FFFFFFF) | (int(stream) << 31)


def box_muller(xa, xb):
    """(radius bits, angle bits) -> two normals, float64 math on the exactly-representable
    fp32 inputs u1 = (2*(xa>>9)+1)*2^-24 in (0,1) and a = (xb>>8)*2^-24 in [0,1)."""
    u1 = (2.0 * (xa >> np.uint32(9)).astype(np.float64) + 1.0) * 2.0 ** -24
    a = (xb >> np.uint32(8)).astype(np.float64) * 2.0 ** -24
    r = np.sqrt(-2.0 * np.log(u1))
    return r * np.cos(2.0 * np.pi * a), r * np.sin(2.0 * np.pi * a)


def box_muller_word(x):
    """Latent-noise stream (stream 0): ONE 32-bit word per pair, 20 radius bits + 12 angle bits, the angle lattice
    dithered by the low 8 radius bits (philox.cuh box_muller_word): u1 = 1 - (x>>12)*2^-20 in (0,1],
    g = 1 + ((x&0xfff)*2^11 + ((x>>12)&0xff)*2^3)*2^-23, theta = fl32(fma(g, fl32(2 pi), fl32(-3 pi)))
    in (-pi,pi).  The device evaluates lg2/sqrt/sin/cos on the special-function unit, so its fp32 normals sit
    within ~1e-6 absolute of this float64 evaluation before rounding to the latent dtype."""
    x = np.asarray(x, dtype=np.uint32)
    u1 = 1.0 - (x >> np.uint32(12)).astype(np.float64) * 2.0 ** -20
    mant = (x & np.uint32(0xFFF)).astype(np.uint64) * np.uint64(2048) + ((x >> np.uint32(12)) & np.uint32(0xFF)).astype(np.uint64) * np.uint64(8)
    g = 1.0 + mant.astype(np.float64) * 2.0 ** -23
    two_pi = np.float64(np.float32(6.283185307179586))
    c = np.float64(np.float32(-9.42477796076938))
    theta = (g * two_pi + c).astype(np.float32).astype(np.float64)
    r = np.sqrt(-2.0 * np.log(u1))
    return r * np.cos(theta), r * np.sin(theta)


def _words(seed, offset, sample, n_calls, stream):
    ctr = np.arange(n_calls, dtype=np.uint64)
    return philox4x32_10(ctr, np.uint64(sample), np.uint64(int(offset) & 0xFFFFFFFF),
                         np.uint64(_counter_hi(offset, stream)),
                         int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF)


def normal_octs(seed, offset, sample, n_octs):
    """Latent normals for octs 0..n_octs-1 of ``sample`` -> float64 array [n_octs*8]: word w of
    call ``oct`` gives elements 8*oct + 2w (cos) and 8*oct + 2w + 1 (sin)."""
    xs = _words(seed, offset, sample, n_octs, 0)
    cols = []
    for x in xs:
        z0, z1 = box_muller_word(x)
        cols += [z0, z1]
    return np.stack(cols, axis=1).reshape(-1)


def normal_quads(seed, offset, sample, n_quads, stream=1):
    """Timestep stream (stream 1): two full words per pair, precise map -> float64 [n_quads*4]."""
    x0, x1, x2, x3 = _words(seed, offset, sample, n_quads, stream)
    z0, z1 = box_muller(x0, x1)
    z2, z3 = box_muller(x2, x3)
    return np.stack([z0, z1, z2, z3], axis=1).reshape(-1)


def noise(seed, offset, batch, chw):
    """fp32 noise [batch, chw] of the generate mode (before rounding to the latent dtype)."""
    n_octs = (chw + 7) // 8
    out = np.empty((batch, chw), dtype=np.float32)
    for b in range(batch):
        out[b] = normal_octs(seed, offset, b, n_octs)[:chw].astype(np.float32)
    return out


def timestep_normals(seed, offset, batch):
    """z [batch] float64 of the timestep stream (u = sigmoid(z))."""
    return np.array([normal_quads(seed, offset, b, 1, stream=1)[0] for b in range(batch)])


def raw_stream(seed, offset, sample, n_quads, stream=0):
    """Raw uint32 words [n_quads,4] — compared bit-exactly against ``yat_philox_raw``."""
    return np.stack(_words(seed, offset, sample, n_quads, stream), axis=1)
