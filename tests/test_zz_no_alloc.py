"""b200pt_trace_queue_static (no counterpart in the reference): the ray-batch query of
b200pt_trace_queue with its scratch in a static __device__ array instead of cudaMalloc'd temporaries. Checked through the C ABI against
the existing entry point, bit for bit (the bar of every hit-id test here), on two calls in a row that share the array. Collected last
(file name) so that a failure here cannot hide another test under `pytest -x`; needs no oracle and no state from other tests."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _rays(rng, n, lo, hi):
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[::13, 2] = -0.0          # negative zero components (the reciprocal's sign decides the slab order)
    d[::17, 0] = 0.0
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.ascontiguousarray(np.concatenate([o, d], axis=1), np.float32)


def test_static_scratch_entry_point_equals_trace_queue(pt, wl, ctx):
    scene = wl.cornell(64, 48)["scene"]
    bvh = ctx.build_bvh(scene.vertices, scene.indices)
    lo, hi = scene.vertices.min(axis=0) - 0.5, scene.vertices.max(axis=0) + 0.5
    rng = np.random.default_rng(2024)
    gold = np.load(os.path.join(GOLDEN, "cornell_64x48_mis.npz"))
    ntri = scene.triangle_count
    # two calls in a row on the SAME static array: a large batch, then a smaller, different one (stale records of call 1 lie behind it)
    batches = []
    n1 = 200_000
    batches.append((_rays(rng, n1, lo, hi), rng.integers(-1, ntri, n1).astype(np.int32)))
    n2 = len(gold["rays"])
    batches.append((np.ascontiguousarray(gold["rays"], np.float32), np.ascontiguousarray(gold["last_prim"], np.int32)))
    assert n2 < n1 <= ctx.STATIC_MAX_RAYS
    cap = n1
    d_prim, d_t, d_uv, d_occ = ctx.malloc(4 * cap), ctx.malloc(4 * cap), ctx.malloc(8 * cap), ctx.malloc(cap)
    try:
        for call, (rays, last) in enumerate(batches):
            n = len(rays)
            t_max = rng.uniform(0.05, 8.0, n).astype(np.float32)
            d_rays, d_last, d_tmax = ctx.upload(rays), ctx.upload(last), ctx.upload(t_max)
            # closest hit
            ctx.trace_queue_static(bvh, d_rays, None, d_last, n, d_prim, d_t, d_uv, None)
            ctx.sync()
            prim, t, uv = ctx.download(d_prim, (n,), np.int32), ctx.download(d_t, (n,), np.float32), ctx.download(d_uv, (n, 2), np.float32)
            prim0, t0, uv0 = ctx.trace_queue(bvh, rays, None, last)
            assert (prim0 >= 0).sum() > n // 10, "the batch should hit geometry"
            assert np.array_equal(prim, prim0), f"call {call}: {int((prim != prim0).sum())} hit ids differ from b200pt_trace_queue"
            assert np.array_equal(t.view(np.uint32), t0.view(np.uint32)) and np.array_equal(uv.view(np.uint32), uv0.view(np.uint32)), f"call {call}: t / uv bits differ"
            # any hit
            ctx.trace_queue_static(bvh, d_rays, d_tmax, d_last, n, None, None, None, d_occ)
            ctx.sync()
            occ = ctx.download(d_occ, (n,), np.uint8)
            occ0 = ctx.trace_queue(bvh, rays, t_max, last)
            assert 0 < int(occ0.sum()) < n
            assert np.array_equal(occ, occ0), f"call {call}: occlusion flags differ"
            if call == 1:       # the committed golden hits of the reference's own render, as in test_golden_ray_batch_hit_ids_bit_exact
                assert (prim == gold["hit_prim"]).all()
            for p in (d_rays, d_last, d_tmax):
                ctx.free(p)
        # sizes beyond the static maximum are refused with a clear message, and nothing is launched
        with pytest.raises(pt.B200ptError, match="B200PT_TRACE_QUEUE_STATIC_MAX_RAYS"):
            ctx.trace_queue_static(bvh, d_prim, None, None, ctx.STATIC_MAX_RAYS + 1, d_prim, d_t, d_uv, None)
        ctx.trace_queue_static(bvh, None, None, None, 0, None, None, None, None)     # empty batch: accepted, no work
        ctx.sync()
    finally:
        for p in (d_prim, d_t, d_uv, d_occ):
            ctx.free(p)
        ctx.destroy_bvh(bvh)
