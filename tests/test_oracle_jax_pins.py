"""External pin of the PRNG oracle: values printed in JAX's documentation (tests/jax_published.py
lists each with its source page).  Rows a9 / n2."""
import numpy as np
import pytest

from oracle import prng
import jax_published as jp   # tests/jax_published.py


@pytest.mark.parametrize("layout,seed,n,want", jp.SPLITS)
def test_split_matches_jax_docs(layout, seed, n, want):
    got = prng.split(jp.key_of(seed), n, layout)
    assert got.dtype == np.uint32 and got.tolist() == want


def _walk(seed, path, layout):
    key = jp.key_of(seed)
    for nsplit, i in path:
        key = prng.split(key, nsplit, layout)[i]
    return key


@pytest.mark.parametrize("layout,seed,path,n,want", jp.NORMALS)
def test_normal_matches_jax_docs(layout, seed, path, n, want):
    got = prng.normal(_walk(seed, path, layout), n, layout)
    np.testing.assert_allclose(got, np.asarray(want, np.float32), rtol=jp.NORMAL_RTOL, atol=0)


@pytest.mark.parametrize("layout,seed,want", jp.UNIFORMS)
def test_uniform_matches_jax_docs(layout, seed, want):
    got = prng.uniform(jp.key_of(seed), 1, layout=layout)
    assert got[0] == np.float32(want)
