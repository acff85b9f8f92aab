"""GPU tests of the multi-GPU split, through the C ABI.  On one GPU: the stride/phase passes of N virtual ranks summed
on the device film equal the single full pass (same samples, so equality up to float summation order).  With >= 2 GPUs
(gpurun --gpus 2): two contexts, sample-index split, ONE ncclReduce of the films (mtsb_comm_create_all +
mtsb_film_reduce_all) against the single-GPU render."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _render(pkg, ctx, sd, spp, world=1, ranks=None, per_pass=0):
    B, M = pkg.binding, pkg.multigpu
    sc = sd.upload(ctx)
    rd = B.Render(sc, sd.camera(), sd.film(), sd.integrator_desc(), seed=77)
    try:
        for r in (range(world) if ranks is None else ranks):
            M.render_distributed(rd, None, spp, world, r, per_pass)
        return rd.film_to_host().astype(np.float64), rd.stats()
    finally:
        rd.close()
        sc.close()


@pytest.mark.parametrize("world,per_pass", [(2, 0), (3, 2), (8, 1)])
def test_virtual_ranks_sum_to_the_single_pass(pkg, ctx, world, per_pass):
    sd = pkg.scenes.materials_scene(40, 24, spp=12, max_depth=5)
    full, st = _render(pkg, ctx, sd, 12)
    split, st2 = _render(pkg, ctx, sd, 12, world=world, per_pass=per_pass)
    assert st["paths"] == st2["paths"] == 40 * 24 * 12
    assert st["closest_rays"] == st2["closest_rays"] and st["shadow_rays"] == st2["shadow_rays"]   # identical paths
    assert np.allclose(split, full, rtol=2e-5, atol=1e-5)
    one, _ = _render(pkg, ctx, sd, 12, world=world, ranks=[0])
    assert one[..., 4].sum() < full[..., 4].sum() * (1.0 / world + 0.1)


def test_two_gpu_nccl_film_reduce(pkg):
    B, M = pkg.binding, pkg.multigpu
    if B.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2); the single-GPU split test above covers the sample arithmetic")
    assert B.lib().mtsb_comm_version() > 20000
    sd = pkg.scenes.materials_scene(64, 40, spp=16, max_depth=5)
    ctxs = [B.Context(0), B.Context(1)]
    comms = B.comm_create_all(ctxs)
    scs = [sd.upload(c) for c in ctxs]
    rds = [B.Render(s, sd.camera(), sd.film(), sd.integrator_desc(), seed=77) for s in scs]
    try:
        for r, rd in enumerate(rds):
            M.render_distributed(rd, None, 16, 2, r)
        B.film_reduce_all(rds, comms, root=0)
        got = rds[0].film_to_host().astype(np.float64)
        other = rds[1].film_to_host().astype(np.float64)
        ref_rd = B.Render(scs[0], sd.camera(), sd.film(), sd.integrator_desc(), seed=77)
        ref_rd.render_pass(0, 16, 16)
        ref = ref_rd.film_to_host().astype(np.float64)
        ref_rd.close()
        assert np.allclose(got, ref, rtol=2e-5, atol=1e-5)
        assert other[..., 4].sum() < 0.6 * ref[..., 4].sum()        # non-root keeps its own share
    finally:
        for x in rds + comms + scs + ctxs:
            x.close()


def test_config3_scale_properties(pkg, ctx):
    """BASELINE config-3 scale (10 485 762 triangles, GGX copper + bk7 glass + diffuse ground, area light + constant
    environment, Gaussian film) at a reduced image size: properties that need no oracle."""
    S, B, M = pkg.scenes, pkg.binding, pkg.multigpu
    sd = S.sphere_field(subdiv=6, nx=16, ny=8, width=240, height=136, spp=4, max_depth=8, rfilter="gaussian")
    sc = sd.upload(ctx, B.BUILDER_PLOC)
    try:
        assert sc.primitive_count == 10485763          # 128 x 81 920 + 2 ground triangles + the light rectangle
        st = sc.build_stats()
        assert st["nodes"] > 1000000 and st["depth"] < 32
        rd = B.Render(sc, sd.camera(), sd.film(), sd.integrator_desc(), seed=5)
        rd.render_pass(0, 4, 4)
        full = rd.film_to_host().astype(np.float64)
        stats = rd.stats()
        rd.clear()
        for r in range(2):
            M.render_distributed(rd, None, 4, 2, r, 1)
        split = rd.film_to_host().astype(np.float64)
        rd.close()
        assert stats["paths"] == 240 * 136 * 4
        assert stats["closest_rays"] >= stats["paths"] and 1.0 <= stats["depth_sum"] / stats["paths"] <= 8.0
        assert np.isfinite(full).all() and (full >= 0).all()
        # every sample deposits filter weights that sum to ~1 (32-bin Gaussian table, normalised as in rfilter.cpp:37-55)
        assert abs(full[..., 4].sum() / (240 * 136 * 4) - 1.0) < 2e-2
        # alpha <= weight everywhere, = weight where every contributing sample hit geometry
        assert (full[..., 3] <= full[..., 4] * (1 + 1e-5) + 1e-6).all()
        assert np.allclose(split, full, rtol=5e-5, atol=1e-5)
        # the ground / spheres / sky all show up: radiance varies, nothing is black overall
        b = sd.film().border
        img = full[b:-b, b:-b, 0:3] / np.maximum(full[b:-b, b:-b, 4:5], 1e-9)
        assert img.mean() > 0.05 and img.std() > 0.01
    finally:
        sc.close()
