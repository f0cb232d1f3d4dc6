"""The reference's trace-time noise constants without jax.

`NoisyMLP.__init__` seeds `jax.random.PRNGKey(1234)` and splits one key per layer
(hanabi_agents/rlax_dqn/noisy_mlp.py:127,140); every `NoisyLinear` wraps its key in an
`hk.PRNGSequence` and draws `jax.random.normal(next(rng), shape)` for the weight noise and then for the
bias noise (:38,79,87) while the network is traced, so the noise is a constant of the compiled
function.  A pickle written by `save_weights` (rlax_rainbow.py:344-357) holds w, b, w_mu, b_mu, w_sigma,
b_sigma but NOT that noise: to evaluate the network the reference evaluates, the constants have to be
regenerated.  This module restates the pieces of jax.random involved (jax is not installable here):

  threefry2x32        the counter-based generator behind every jax key (Random123, 20 rounds)
  split               `jax.random.split`: threefry over iota(2 * num), reshaped to [num, 2]
  random_bits         `_random_bits(key, 32, shape)`: threefry over iota(size), padded to even length
  uniform / normal    mantissa-fill uniform in [nextafter(-1, 0), 1) and sqrt(2) * erf_inv(u), float32

Pinned on the known answers jax itself tests and documents (tests/test_jax_prng.py): the three Random123
vectors of `random_test.py::testThreefry2x32`, `split(PRNGKey(0))` and `normal(PRNGKey(0), (1,))` from the
jax documentation (erf_inv is XLA's float32 polynomial, restated: the documented draws come out to the
last digit; a fused multiply-add inside XLA's Horner steps could still move the last bit of some values,
far below the 1e-3 tolerance of the Q-values).  `hk.PRNGSequence.__next__` is `key, sub = split(key); return sub` in every
haiku release (with or without its reserve queue).
"""
import numpy as np

_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))


def _rotl(x, d):
    return (x << np.uint32(d)) | (x >> np.uint32(32 - d))


def threefry2x32(key, x0, x1):
    """Threefry-2x32, 20 rounds.  key: two uint32; x0, x1: uint32 arrays of equal shape."""
    k0, k1 = np.uint32(key[0]), np.uint32(key[1])
    ks = (k0, k1, k0 ^ k1 ^ np.uint32(0x1BD11BDA))
    x0 = np.array(x0, dtype=np.uint32, copy=True)
    x1 = np.array(x1, dtype=np.uint32, copy=True)
    with np.errstate(over="ignore"):
        x0 += ks[0]
        x1 += ks[1]
        for i in range(5):
            for r in _ROT[i % 2]:
                x0 += x1
                x1 = _rotl(x1, r)
                x1 ^= x0
            x0 += ks[(i + 1) % 3]
            x1 += ks[(i + 2) % 3] + np.uint32(i + 1)
    return x0, x1


def _threefry_counts(key, counts):
    """jax.random.threefry_2x32(key, counts): the count array is cut into two halves that form the
    (x0, x1) pairs; an odd length is padded with one zero and the last output dropped."""
    counts = np.asarray(counts, np.uint32).ravel()
    odd = counts.size % 2
    if odd:
        counts = np.concatenate([counts, np.zeros(1, np.uint32)])
    h = counts.size // 2
    y0, y1 = threefry2x32(key, counts[:h], counts[h:])
    out = np.concatenate([y0, y1])
    return out[:-1] if odd else out


def prng_key(seed):
    """jax.random.PRNGKey(seed) for a non-negative 32-bit seed (x64 disabled): [0, seed]."""
    return np.array([(int(seed) >> 32) & 0xFFFFFFFF, int(seed) & 0xFFFFFFFF], np.uint32)


def split(key, num=2):
    return _threefry_counts(key, np.arange(2 * num, dtype=np.uint32)).reshape(num, 2)


def random_bits(key, shape):
    size = int(np.prod(shape)) if len(shape) else 1
    return _threefry_counts(key, np.arange(size, dtype=np.uint32)).reshape(shape)


def uniform(key, shape, minval=0.0, maxval=1.0):
    bits = random_bits(key, shape)
    f = ((bits >> np.uint32(9)) | np.uint32(0x3F800000)).view(np.float32) - np.float32(1.0)
    lo, hi = np.float32(minval), np.float32(maxval)
    return np.maximum(lo, f * (hi - lo) + lo).astype(np.float32)


# XLA's float32 ErfInv (xla/client/lib/math.cc): Giles' single-precision polynomials in w = -log((1-x)(1+x))
_ERFINV_LT5 = (2.81022636e-08, 3.43273939e-07, -3.5233877e-06, -4.39150654e-06, 0.00021858087,
               -0.00125372503, -0.00417768164, 0.246640727, 1.50140941)
_ERFINV_GE5 = (-0.000200214257, 0.000100950558, 0.00134934322, -0.00367342844, 0.00573950773,
               -0.0076224613, 0.00943887047, 1.00167406, 2.83297682)


def erf_inv32(x):
    f = np.float32
    x = np.asarray(x, f)
    with np.errstate(divide="ignore"):
        w = (-np.log((f(1) - x) * (f(1) + x))).astype(f)
    lt = w < f(5)
    ww = np.where(lt, w - f(2.5), np.sqrt(np.maximum(w, f(0))) - f(3)).astype(f)
    p = np.where(lt, f(_ERFINV_LT5[0]), f(_ERFINV_GE5[0])).astype(f)
    for a, b in zip(_ERFINV_LT5[1:], _ERFINV_GE5[1:]):
        p = (np.where(lt, f(a), f(b)) + p * ww).astype(f)
    return np.where(np.abs(x) == f(1), x * f(np.inf), p * x).astype(f)


def normal(key, shape):
    lo = np.nextafter(np.float32(-1.0), np.float32(0.0))
    u = uniform(key, shape, lo, 1.0)
    return (np.float32(np.sqrt(2.0)) * erf_inv32(u)).astype(np.float32)


def noisy_mlp_noise(sizes, seed=1234):
    """[(eps_w [in, out], eps_b [out])] per layer of NoisyMLP(sizes[1:]) applied to sizes[0] inputs
    (noisy_mlp.py:127,140 + :38,79,87)."""
    rng = prng_key(seed)
    out = []
    for fi, fo in zip(sizes[:-1], sizes[1:]):
        key, rng = split(rng)                  # key, self.rng = jax.random.split(self.rng)
        seq, sub = split(key)                  # next(hk.PRNGSequence(key))
        eps_w = normal(sub, (fi, fo))
        seq, sub = split(seq)
        eps_b = normal(sub, (fo,))
        out.append((eps_w, eps_b))
    return out
