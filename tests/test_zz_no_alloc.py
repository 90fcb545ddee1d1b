"""GPU: pgi_rank_batch_noalloc (static __device__ scratch, no device allocation) against pgi_rank_batch, bit-exact,
through the C ABI, on two calls in a row (the second call finds the first call's contents in the shared scratch array).
Collected last on purpose (file name)."""
import ctypes
import numpy as np, pytest
import datasets
from conftest import load_package

pytestmark = pytest.mark.gpu

@pytest.fixture(scope="module")
def capi():
    c = load_package().capi
    assert c.lib().pgi_device_count() > 0, "no CUDA device visible: the product has no CPU path"
    return c

def cases():
    s = datasets.synth_index(encoded=True)
    return {"xy_legacy": datasets.gold("xy.ri"), "synth_encoded": s["ri"]}

@pytest.mark.parametrize("name", ["xy_legacy", "synth_encoded"])
def test_rank_batch_noalloc_equals_rank_batch_twice(capi, name):
    g = capi.Index(cases()[name]); n = g.info["n"]
    rng = np.random.default_rng(11)
    first = np.concatenate([rng.integers(0, n + 1, capi.RANK_NOALLOC_MAX - 4), [0, n, n - 1, 1]]).astype(np.uint64)   # the maximum size
    second = rng.integers(0, 2 * n, 3001).astype(np.uint64)        # shorter, other positions (some beyond n: read as n), stale scratch behind it
    for pos in (first, second):
        want = g.rank_batch(pos)
        got = g.rank_batch_noalloc(pos)
        assert got.shape == want.shape and np.array_equal(got, want)
    assert np.array_equal(g.rank_batch_noalloc(first), g.rank_batch(first))                  # and the long one again after the short one
    assert g.rank_batch_noalloc(np.zeros(0, dtype=np.uint64)).shape == (0, 8)

def test_rank_batch_noalloc_rejects_sizes_beyond_the_static_scratch(capi):
    g = capi.Index(cases()["xy_legacy"])
    pos = np.zeros(capi.RANK_NOALLOC_MAX + 1, dtype=np.uint64); out = np.zeros((len(pos), 8), dtype=np.uint64)
    L = capi.lib()
    rc = L.pgi_rank_batch_noalloc(g.h, len(pos), pos.ctypes.data, out.ctypes.data)
    assert rc == -1                                                  # PGI_EINVAL
    msg = L.pgi_last_error().decode()
    assert "PGI_RANK_NOALLOC_MAX" in msg and str(capi.RANK_NOALLOC_MAX) in msg
    assert not out.any()                                             # nothing was written
    with pytest.raises(capi.PgiError):
        g.rank_batch_noalloc(pos)
