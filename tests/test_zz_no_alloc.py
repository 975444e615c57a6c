"""bricks_copy_regions_static: the region copies through the entry point that allocates nothing (its
link table lives in a file-scope static __device__ array of the library).  No counterpart in the
reference.  Checked against the existing entry point
(bricks_links_upload + bricks_copy_regions), byte for byte -- the bar these tests use for byte work --
on two calls in a row that share the static table.  Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _links(rng, d_src, d_dst, n_elems, sizes16):
    """ragged links (sizes in 16-byte units) from random places of src to packed places of dst"""
    links, pos = [], 0
    for nb16 in sizes16:
        ne = nb16 * 2  # int64 elements
        s0 = int(rng.integers(0, n_elems - ne) // 2 * 2)
        links.append((d_src + s0 * 8, d_dst + pos * 8, ne * 8))
        pos += ne + 2 * int(rng.integers(0, 5))
    assert pos <= n_elems
    return links


def test_static_copy_matches_the_allocating_entry_point_twice_in_a_row(ctx, built):
    rng = np.random.default_rng(11)
    n = 1 << 20
    src = rng.integers(1, 2 ** 62, n, dtype=np.int64)
    d_src = ctx.upload(src)
    d_new = ctx.upload(np.zeros(n, np.int64))
    d_old = ctx.upload(np.zeros(n, np.int64))
    # first call: many links, ragged sizes, one longer than a 64 KiB chunk; second call: FEWER links, so
    # the shared table still holds stale entries of the first call past n
    for sizes in ([1, 2, 3, 255, 256, 257, 4096, 4097, 70000, 1, 9000] + [5] * 300, [7, 4100, 1, 33]):
        ctx.memset(d_new, 0, n * 8)
        ctx.memset(d_old, 0, n * 8)
        l_new = _links(np.random.default_rng(len(sizes)), d_src, d_new, n, sizes)
        l_old = [(s, d - d_new + d_old, b) for s, d, b in l_new]
        built.copy_regions_static(ctx, l_new)   # the new entry point, through the C ABI
        built.copy_regions(ctx, l_old)          # the existing one
        ctx.sync()
        got = ctx.download(d_new, n, np.int64)
        exp = ctx.download(d_old, n, np.int64)
        assert exp.any()
        assert np.array_equal(got, exp)
    for p in (d_src, d_new, d_old):
        ctx.free(p)


def test_static_copy_rejects_tables_beyond_its_maximum(ctx, built):
    d = ctx.upload(np.zeros(64, np.int64))
    too_many = [(d, d + 256, 16)] * (built.COPY_STATIC_MAX_LINKS + 1)
    with pytest.raises(built.BricksError, match="at most %d" % built.COPY_STATIC_MAX_LINKS):
        built.copy_regions_static(ctx, too_many)
    built.copy_regions_static(ctx, [])  # nothing to do is not an error
    ctx.sync()
    ctx.free(d)
