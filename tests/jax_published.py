"""Values that JAX's own documentation prints for jax.random, used as the EXTERNAL pin of the PRNG
oracle (oracle/prng.py) and of the device PRNG (csrc/prng.cu).  Nothing here was produced by this
repository: each number is a literal output shown in the JAX docs, so rows a9 / n2 of SURVEY.md 8 are
pinned against JAX itself although JAX cannot run in this image.  Integer outputs (key data) are
compared bit-exactly; normals to 2e-6 relative (the docs print 8 significant digits and XLA's erf_inv
polynomial moved by an ulp or two between releases).

`key(s)` / `PRNGKey(s)` = uint32[2] (0, s) for the default threefry2x32 implementation.

LEGACY stream layout (jax_threefry_partitionable = False, the default before JAX 0.5.0):
  [QS]  docs/notebooks/quickstart ("JAX Quickstart"):
          key = random.PRNGKey(0); x = random.normal(key, (10,)); print(x)
          -> [-0.3721109 0.26423115 -0.18252768 -0.7368197 -0.44030377 -0.1521442 -0.67135346 -0.5908641
              0.73168886 0.5673026]
  [SB]  docs/notebooks/Common_Gotchas_in_JAX ("JAX - The Sharp Bits", "Random numbers"):
          key = random.PRNGKey(0);  random.normal(key, shape=(1,))              -> [-0.20584226]
          key, subkey = random.split(key); random.normal(subkey, shape=(1,))    -> [-1.2515389]
  [RN]  docs/jax-101/05-random-numbers ("Pseudo Random Numbers in JAX"):
          key = random.PRNGKey(42);  random.normal(key)                          -> -0.18471177
          new_key, subkey = random.split(key); random.normal(subkey)             -> 1.3694694
          subkeys = random.split(random.PRNGKey(42), 3); normal per subkey       -> [-0.04838839 0.10796146 -1.2226542]
          random.normal(random.PRNGKey(42), shape=(3,))                          -> [0.18693541 -1.2806507 -1.5593133]
  [API] jax.random module documentation / JEP 9263 examples:
          random.split(random.PRNGKey(0)) -> [[4146024105 967050713] [2718843009 1272950319]]
          random.uniform(random.PRNGKey(0)) -> 0.41845703
PARTITIONABLE layout (jax_threefry_partitionable = True, the default since JAX 0.5.0):
  [SB5] the same Sharp Bits section, JAX >= 0.5:
          key = random.key(0);  random.normal(key, shape=(1,))                   -> [1.6226422]
          key, subkey = random.split(key); random.normal(subkey, shape=(1,))     -> [-2.4424558]
  [RN5] docs/random-numbers.md, JAX >= 0.5:
          random.normal(random.key(42))                                          -> -0.028304616
          random.key_data(random.split(random.key(0))) -> [[1797259609 2579123966] [928981903 3453687069]]
          random.normal(random.key(0), shape=(3,))                               -> [1.6226422 2.0252647 -0.43359444]
"""
import numpy as np

LEGACY, PARTITIONABLE = 0, 1

# (layout, seed, n, key data of split(key(seed), n))
SPLITS = [
    (LEGACY, 0, 2, [[4146024105, 967050713], [2718843009, 1272950319]]),
    (PARTITIONABLE, 0, 2, [[1797259609, 2579123966], [928981903, 3453687069]]),
]

# (layout, seed, path, n, values): follow `path` (index into split(key, nsplit) at every hop, given as
# (nsplit, index) pairs), then draw n normals from the key reached.
NORMALS = [
    (LEGACY, 0, (), 10, [-0.3721109, 0.26423115, -0.18252768, -0.7368197, -0.44030377, -0.1521442, -0.67135346,
                         -0.5908641, 0.73168886, 0.5673026]),
    (LEGACY, 0, (), 1, [-0.20584226]),
    (LEGACY, 0, ((2, 1),), 1, [-1.2515389]),
    (LEGACY, 42, (), 1, [-0.18471177]),
    (LEGACY, 42, ((2, 1),), 1, [1.3694694]),
    (LEGACY, 42, ((3, 0),), 1, [-0.04838839]),
    (LEGACY, 42, ((3, 1),), 1, [0.10796146]),
    (LEGACY, 42, ((3, 2),), 1, [-1.2226542]),
    (LEGACY, 42, (), 3, [0.18693541, -1.2806507, -1.5593133]),
    (PARTITIONABLE, 0, (), 1, [1.6226422]),
    (PARTITIONABLE, 0, ((2, 1),), 1, [-2.4424558]),
    (PARTITIONABLE, 42, (), 1, [-0.028304616]),
    (PARTITIONABLE, 0, (), 3, [1.6226422, 2.0252647, -0.43359444]),
]
NORMAL_RTOL = 2e-6

# (layout, seed, value): uniform(key(seed)) in [0, 1) -- exact: a 23-bit mantissa fill, no polynomial
UNIFORMS = [
    (LEGACY, 0, 0.41845703),
]


def key_of(seed):
    return np.array([0, seed], np.uint32)
