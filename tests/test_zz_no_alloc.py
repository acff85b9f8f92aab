"""mtsb_trace_closest_noalloc — the closest-hit query that allocates nothing (include/mtsb200.h; no counterpart in the
reference) against the ordinary entry point mtsb_trace_closest, through the
C ABI.  Integer / record work: the two must agree BIT FOR BIT (t, u, v, prim), which is tighter than the tests' t tolerance.

The file name sorts last on purpose: pytest collects it after every other test, so the static device scratch is exercised
both in a fresh process (`pytest tests/test_zz_no_alloc.py`) and after the rest of the suite has used the device."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scene(pkg, ctx, curved=False):
    S = pkg.scenes
    sd = S.SceneDesc()
    V, N, F = S.displaced_sphere(4, seed=11)
    sd.add_mesh(V, N, F, bsdf=dict(type="diffuse"))
    if curved:      # analytic shapes select the other kernel instantiation (CURVED)
        sd.add_sphere((0.0, 0.0, 2.5), 0.75, bsdf=dict(type="diffuse"))
    return sd.upload(ctx, pkg.binding.BUILDER_PLOC)


def _same(a, b):
    return a.tobytes() == b.tobytes()


@pytest.mark.parametrize("curved", [False, True])
def test_noalloc_matches_trace_closest_two_calls_in_a_row(pkg, ctx, curved):
    """Host pointers.  Two calls in a row with different batches, the second one smaller: whatever the first call left in
    the shared scratch (work counter at n, staged rays, hit records) must not leak into the second."""
    S, B = pkg.scenes, pkg.binding
    sc = _scene(pkg, ctx, curved)
    try:
        first = S.chord_rays(B.NOALLOC_MAX_RAYS, radius=1.6 if not curved else 3.5, seed=101)      # exactly the limit
        second = S.chord_rays(12345, radius=1.2 if not curved else 3.5, seed=202, mint=S.EPSILON)
        for rays in (first, second):
            ref = sc.ray_intersect(rays)
            got = sc.ray_intersect_noalloc(rays)
            assert got.dtype == B.hit_dtype and got.shape == ref.shape
            hit = ref["prim"] != B.MTSB_NO_HIT
            assert 0.05 < hit.mean() < 1.0                     # the batch exercises hits and misses
            assert (got["prim"] == ref["prim"]).all()
            np.testing.assert_allclose(got["t"][hit], ref["t"][hit], rtol=1e-5, atol=0)      # the tests' tolerance ...
            assert _same(got, ref)                                                        # ... and in fact bit for bit
        # and once more with the first batch: same answer as before the smaller call
        assert _same(sc.ray_intersect_noalloc(first), sc.ray_intersect(first))
    finally:
        sc.close()


def test_noalloc_device_pointers(pkg, ctx):
    """on_device = 1: the caller's device-addressable buffers are read and written in place, only the work counter is scratch.
    The buffers are page-locked host memory (mtsb_host_alloc), which the device addresses directly under unified addressing —
    what the per-ray overlay does with its small batches; no other allocator is involved."""
    S, B = pkg.scenes, pkg.binding
    sc = _scene(pkg, ctx)
    bufs = []
    try:
        for n, seed in ((40000, 5), (777, 6)):
            rays = S.chord_rays(n, radius=1.3, seed=seed)
            d_rays, d_ref, d_got = B.PinnedArray((n,), B.ray_dtype), B.PinnedArray((n,), B.hit_dtype), B.PinnedArray((n,), B.hit_dtype)
            bufs += [d_rays, d_ref, d_got]
            d_rays.array[:] = rays
            d_ref.array.view(np.uint8)[:] = 0
            d_got.array.view(np.uint8)[:] = 0xEE
            sc.trace_device("closest", d_rays.ptr, n, d_ref.ptr)
            sc.trace_device_noalloc(d_rays.ptr, n, d_got.ptr)
            ctx.sync()
            assert _same(d_got.array, d_ref.array)
            assert _same(d_ref.array, sc.ray_intersect(rays))
    finally:
        for b in bufs:
            b.free()
        sc.close()


def test_noalloc_rejects_batches_beyond_the_scratch(pkg, ctx):
    import re
    S, B = pkg.scenes, pkg.binding
    assert int(re.search(r"#define\s+MTSB_NOALLOC_MAX_RAYS\s+(\d+)u", open(B.HEADER_PATH).read()).group(1)) == B.NOALLOC_MAX_RAYS
    sc = _scene(pkg, ctx)
    try:
        rays = S.chord_rays(B.NOALLOC_MAX_RAYS + 1, radius=1.3, seed=9)
        with pytest.raises(B.MtsbError) as e:
            sc.ray_intersect_noalloc(rays)
        assert e.value.code == -1 and "65536" in str(e.value)           # MTSB_EINVAL, the message names the limit
        assert len(sc.ray_intersect_noalloc(rays[:0])) == 0              # empty batch: nothing to do
        # the refusal left the entry point usable
        assert _same(sc.ray_intersect_noalloc(rays[:1000]), sc.ray_intersect(rays[:1000]))
    finally:
        sc.close()
