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
  stream 0 = latent noise (quad = element_index // 4 inside the sample, 4 normals per call),
  stream 1 = timestep draw (quad = 0, first normal of the call).
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


def box_muller_fast(xa, xb):
    """Latent-noise stream (stream 0), float64 evaluation of the device map (philox.cuh
    box_muller_fast): u1 = 1 - (xa>>9)*2^-23 in (0,1], theta = 2*pi*((xb>>9)*2^-23 - 0.5) in
    [-pi,pi).  The device evaluates lg2/sqrt/sin/cos on the special-function unit, so its fp32
    normals sit within ~1e-6 absolute of these before rounding to the latent dtype."""
    u1 = 1.0 - (xa >> np.uint32(9)).astype(np.float64) * 2.0 ** -23
    theta = 2.0 * np.pi * ((xb >> np.uint32(9)).astype(np.float64) * 2.0 ** -23 - 0.5)
    r = np.sqrt(-2.0 * np.log(u1))
    return r * np.cos(theta), r * np.sin(theta)


def normal_quads(seed, offset, sample, n_quads, stream=0):
    """Normals for quads 0..n_quads-1 of ``sample`` -> float64 array [n_quads*4].  Stream 0
    (latent noise) uses the fast map, stream 1 (timestep draw) the precise one."""
    quad = np.arange(n_quads, dtype=np.uint64)
    x0, x1, x2, x3 = philox4x32_10(quad, np.uint64(sample), np.uint64(int(offset) & 0xFFFFFFFF),
                                   np.uint64(_counter_hi(offset, stream)),
                                   int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF)
    bm = box_muller_fast if stream == 0 else box_muller
    z0, z1 = bm(x0, x1)
    z2, z3 = bm(x2, x3)
    return np.stack([z0, z1, z2, z3], axis=1).reshape(-1)


def noise(seed, offset, batch, chw):
    """fp32 noise [batch, chw] of the generate mode (before rounding to the latent dtype)."""
    n_quads = (chw + 3) // 4
    out = np.empty((batch, chw), dtype=np.float32)
    for b in range(batch):
        out[b] = normal_quads(seed, offset, b, n_quads, stream=0)[:chw].astype(np.float32)
    return out


def timestep_normals(seed, offset, batch):
    """z [batch] float64 of the timestep stream (u = sigmoid(z))."""
    return np.array([normal_quads(seed, offset, b, 1, stream=1)[0] for b in range(batch)])


def raw_stream(seed, offset, sample, n_quads, stream=0):
    """Raw uint32 words [n_quads,4] — compared bit-exactly against ``yat_philox_raw``."""
    quad = np.arange(n_quads, dtype=np.uint64)
    xs = philox4x32_10(quad, np.uint64(sample), np.uint64(int(offset) & 0xFFFFFFFF),
                       np.uint64(_counter_hi(offset, stream)),
                       int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF)
    return np.stack(xs, axis=1)
