"""Oracle: threefry2x32 counter PRNG with jax.random's split / bits / uniform / normal.

Restates jax/_src/prng.py (threefry_2x32, threefry_split, threefry_random_bits) and
jax/_src/random.py (_uniform, _normal_real) from the published algorithm -- jax is an
unpinned dependency of the reference (requirements.txt:2,9) and is absent here.
Keys enter the reference at src/engine_run.py:42 and experiments/uci_benchmark.py:363,390.
Known-answer vectors are the Random123 threefry2x32-20 KATs (tests/test_oracle_prng.py).

Two stream layouts exist because ``jax_threefry_partitionable`` flipped default in
JAX 0.5.0 and the reference leaves JAX unpinned:
  LAYOUT_LEGACY        counters = iota(n) cut in two halves, outputs concatenated
  LAYOUT_PARTITIONABLE counter  = (hi, lo) 64-bit flat index per element; bits = out0 ^ out1

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Parity unpinned.
"""

import numpy as np

LAYOUT_LEGACY = 0
LAYOUT_PARTITIONABLE = 1

_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))


def _rotl(x, r):
    return ((x << np.uint32(r)) | (x >> np.uint32(32 - r))).astype(np.uint32)


def threefry2x32(key, x0, x1):
    """20-round Threefry-2x32.  key = (k0, k1) uint32; x0, x1 uint32 arrays."""
    with np.errstate(over="ignore"):
        k0, k1 = np.uint32(key[0]), np.uint32(key[1])
        ks = (k0, k1, np.uint32(k0 ^ k1 ^ np.uint32(0x1BD11BDA)))
        x0 = (np.asarray(x0, np.uint32) + ks[0]).astype(np.uint32)
        x1 = (np.asarray(x1, np.uint32) + ks[1]).astype(np.uint32)
        for g in range(5):
            for r in _ROT[g % 2]:
                x0 = (x0 + x1).astype(np.uint32)
                x1 = _rotl(x1, r) ^ x0
            x0 = (x0 + ks[(g + 1) % 3]).astype(np.uint32)
            x1 = (x1 + ks[(g + 2) % 3] + np.uint32(g + 1)).astype(np.uint32)
    return x0, x1


def _bits_legacy(key, n):
    cnt = np.arange(n, dtype=np.uint32)
    if n % 2:
        cnt = np.concatenate([cnt, np.zeros(1, np.uint32)])
    half = cnt.size // 2
    o0, o1 = threefry2x32(key, cnt[:half], cnt[half:])
    return np.concatenate([o0, o1])[:n]


def random_bits(key, n, layout=LAYOUT_PARTITIONABLE):
    """n uint32 words for a flat shape (n,)."""
    if layout == LAYOUT_LEGACY:
        return _bits_legacy(key, n)
    idx = np.arange(n, dtype=np.uint64)
    hi = (idx >> np.uint64(32)).astype(np.uint32)
    lo = (idx & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    o0, o1 = threefry2x32(key, hi, lo)
    return o0 ^ o1


def split(key, num=2, layout=LAYOUT_PARTITIONABLE):
    """jax.random.split -> (num, 2) uint32."""
    if layout == LAYOUT_LEGACY:
        return _bits_legacy(key, 2 * num).reshape(num, 2)
    o0, o1 = threefry2x32(key, np.zeros(num, np.uint32), np.arange(num, dtype=np.uint32))
    return np.stack([o0, o1], axis=-1)


def fold_in(key, data):
    """jax.random.fold_in: threefry(key, seed-key-of(data)) where seed key = (0, data)."""
    o0, o1 = threefry2x32(key, np.zeros(1, np.uint32), np.asarray([data], np.uint32))
    return np.array([o0[0], o1[0]], np.uint32)


def bits_to_unit_float(bits):
    """23 mantissa bits -> float32 in [0, 1)."""
    f = ((bits >> np.uint32(9)) | np.uint32(0x3F800000)).view(np.float32)
    return f - np.float32(1.0)


def uniform(key, n, minval=0.0, maxval=1.0, layout=LAYOUT_PARTITIONABLE):
    u = bits_to_unit_float(random_bits(key, n, layout))
    lo, hi = np.float32(minval), np.float32(maxval)
    return np.maximum(lo, u * (hi - lo) + lo).astype(np.float32)


def erfinv32(x):
    """Giles' single-precision erfinv polynomial (the one XLA lowers erf_inv f32 to)."""
    x = np.asarray(x, np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        w = -np.log1p(-(x * x)).astype(np.float32)
    wa = w - np.float32(2.5)
    pa = np.float32(2.81022636e-08)
    for c in (3.43273939e-07, -3.5233877e-06, -4.39150654e-06, 0.00021858087,
              -0.00125372503, -0.00417768164, 0.246640727, 1.50140941):
        pa = np.float32(c) + pa * wa
    wb = np.sqrt(np.maximum(w, 0)).astype(np.float32) - np.float32(3.0)
    pb = np.float32(-0.000200214257)
    for c in (0.000100950558, 0.00134934322, -0.00367342844, 0.00573950773,
              -0.0076224613, 0.00943887047, 1.00167406, 2.83297682):
        pb = np.float32(c) + pb * wb
    p = np.where(w < np.float32(5.0), pa, pb).astype(np.float32)
    return (p * x).astype(np.float32)


def normal_from_bits(bits):
    """jax.random.normal's map from raw 32-bit words to N(0,1) float32."""
    lo = np.nextafter(np.float32(-1.0), np.float32(0.0))
    hi = np.float32(1.0)
    u = np.maximum(lo, bits_to_unit_float(bits) * (hi - lo) + lo).astype(np.float32)
    return (np.float32(np.sqrt(2.0)) * erfinv32(u)).astype(np.float32)


def normal(key, n, layout=LAYOUT_PARTITIONABLE):
    return normal_from_bits(random_bits(key, n, layout))


def randint_jax(key, span, layout=LAYOUT_PARTITIONABLE):
    """jax.random.randint(key, (), 0, span) for int32 (jax/_src/random.py::_randint): two words from
    split(key); ((hi % span) * (2^32 % span) + lo % span) % span in wrapping uint32 arithmetic."""
    k = split(key, 2, layout)
    hi = np.uint32(random_bits(k[0], 1, layout)[0])
    lo = np.uint32(random_bits(k[1], 1, layout)[0])
    sp = np.uint32(span)
    with np.errstate(over="ignore"):
        mult = np.uint32(65536) % sp
        mult = np.uint32(mult * mult) % sp
        off = np.uint32(np.uint32(np.uint32(hi % sp) * mult) + np.uint32(lo % sp)) % sp
    return int(off)


def randint_bits_to_index(bits, span):
    """Uniform integer in [0, span) from one 32-bit word by multiply-shift (device rule)."""
    return ((bits.astype(np.uint64) * np.uint64(span)) >> np.uint64(32)).astype(np.int32)
