"""GPU parity tests proper: the CUDA path, called through the C ABI, against the oracle (the reference's own CPU kernels)
and against the committed golden vectors. Bars (BASELINE.json north_star): hit primitive ids bit-exact on fixed ray
batches (exact distance ties excluded); radiance within 1e-3 mean relative error with matched RNG streams."""
import importlib
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _rel_err(ref_img, gpu_img):
    return float((np.abs(ref_img - gpu_img) / np.maximum(ref_img, 1e-3)).mean())


@pytest.fixture(scope="module")
def cornell_gpu(pt, wl, ctx):
    cfg = wl.cornell(64, 48)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    yield cfg, g
    g.close()


def test_golden_ray_batch_hit_ids_bit_exact(pt, ctx, cornell_gpu):
    """No oracle needed: committed rays + hits logged from the reference's own render (tools/make_golden.py)."""
    cfg, g = cornell_gpu
    gold = np.load(os.path.join(GOLDEN, "cornell_64x48_mis.npz"))
    prim, t, uv = ctx.trace_closest(g.bvh, gold["rays"], gold["last_prim"])
    assert (prim == gold["hit_prim"]).all()
    hit = prim >= 0
    assert np.array_equal(t[hit], gold["hit_t"][hit]) and np.array_equal(uv[hit], gold["hit_uv"][hit])   # same bits, not just same ids


@pytest.mark.parametrize("variant,lss", [("mis", "mis"), ("ris", "ris")])
def test_golden_image(pt, wl, ctx, variant, lss):
    gold = np.load(os.path.join(GOLDEN, f"cornell_64x48_{variant}.npz"))
    cfg = wl.cornell(int(gold["width"]), int(gold["height"]))
    cfg["options"] = dict(lss=lss, ess="alias_table" if variant == "ris" else "none", ggx=0, later=3)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    g.render_frame()
    g.synchronize()
    assert (g.download("first_hit_prim_index") == gold["first_hit_prim_index"]).all()
    assert np.array_equal(g.download("shading_normals"), gold["shading_normals"])
    assert _rel_err(gold["pixels_1spp"], g.download("pixels")) < 1e-3
    for _ in range(int(gold["spp"]) - 1):
        g.render_frame()
    g.synchronize()
    spp = int(gold["spp"])
    assert _rel_err(gold["pixels"] / spp, g.download("pixels") / spp) < 1e-3
    g.close()


def test_oracle_ray_batches_closest_and_shadow(pt, wl, ctx, ref):
    """Fresh ray batches: primary + bounce + shadow rays logged from an oracle render at 256x192, plus random shadow queries."""
    cfg = wl.cornell(256, 192)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    r.ray_log_begin(2_000_000, 1)
    r.render(2)
    log = r.ray_log_end(2_000_000)
    assert len(log["rays"]) > 300_000
    prim, t, _ = ctx.trace_closest(g.bvh, log["rays"], log["last_prim"])
    ties = (prim != log["hit_prim"]) & (t == log["t"])
    assert ((prim == log["hit_prim"]) | ties).all(), f"{int((prim != log['hit_prim']).sum())} hit-id mismatches"
    same = prim == log["hit_prim"]
    assert np.array_equal(t[same], log["t"][same])
    # shadow rays: occlusion predicate of evaluate_shadow_ray (Intersect.h:224-286) on the same rays with random limits
    rng = np.random.default_rng(3)
    n = 200_000
    rays, last = log["rays"][:n], log["last_prim"][:n]
    t_max = np.where(log["t"][:n] > 0, log["t"][:n] * rng.uniform(0.5, 1.5, n).astype(np.float32), rng.uniform(0.1, 10.0, n).astype(np.float32)).astype(np.float32)
    occ_gpu = ctx.trace_shadow(g.bvh, rays, t_max, last)
    occ_ref = r.trace_shadow(rays, t_max, last)
    assert (occ_gpu == occ_ref).all()
    g.close()


def _assert_hits_equal(prim, t, log, what):
    ties = (prim != log["hit_prim"]) & (t == log["t"])
    bad = ~((prim == log["hit_prim"]) | ties)
    assert not bad.any(), f"{what}: {int(bad.sum())} hit-id mismatches of {len(prim)} rays ({int(ties.sum())} exact-distance ties excluded)"
    same = prim == log["hit_prim"]
    assert np.array_equal(t[same], log["t"][same]), f"{what}: hit distances differ in their bits"


def test_trace_queue_flavours_agree(pt, wl, ctx, ref):
    """The ray-queue traversal kernel (the kernel FullPathTracer traces with) in each loop flavour -- classic, pipelined at 5 / 6 / 7
    resident blocks per SM -- against the oracle's logged rays: same ids, same bits of t and uv as the per-thread traversal."""
    cfg = wl.cornell(256, 192)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    r.ray_log_begin(1_000_000, 1)
    r.render(1)
    log = r.ray_log_end(1_000_000)
    prim0, t0, uv0 = ctx.trace_closest(g.bvh, log["rays"], log["last_prim"])
    rng = np.random.default_rng(5)
    n = len(log["rays"])
    t_max = np.where(log["t"] > 0, log["t"] * rng.uniform(0.5, 1.5, n).astype(np.float32), rng.uniform(0.1, 10.0, n).astype(np.float32)).astype(np.float32)
    occ0 = ctx.trace_shadow(g.bvh, log["rays"], t_max, log["last_prim"])
    ctx.counters_enable(True)
    ctx.counters_read(True)
    try:
        for flavour in (0, 2, 5, 6, 7):
            ctx.set_trace_flavour(flavour)
            prim, t, uv = ctx.trace_queue(g.bvh, log["rays"], None, log["last_prim"])
            _assert_hits_equal(prim, t, log, f"flavour {flavour}")
            assert np.array_equal(prim, prim0) and np.array_equal(t, t0) and np.array_equal(uv, uv0), f"flavour {flavour} differs from the per-thread traversal"
            occ = ctx.trace_queue(g.bvh, log["rays"], t_max, log["last_prim"])
            assert np.array_equal(occ, occ0), f"flavour {flavour}: occlusion flags differ"
        c = ctx.counters_read(True)
        assert c["stack_overflows"] == 0 and c["queue_rays"] == 10 * n
    finally:
        ctx.set_trace_flavour(-1)
        ctx.counters_enable(False)
    g.close()


def test_c3_hit_ids_on_the_10m_triangle_bvh(pt, wl, ctx, ref):
    """SURVEY 8(d) C3 at its stated size: the 9 998 802-triangle scene's own BVH8 (about a million nodes: the deepest tree, the stack
    spill path and the SAH collapse are only stressed here). >= 2e6 primary + bounce + shadow rays logged from the reference's CPU
    render of that scene are traced through the per-thread traversal and through the ray-queue kernel: ids and t bit-exact,
    occlusion predicates equal, no traversal-stack overflow."""
    cfg = wl.synthetic(480, 270)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    bvh = ctx.build_bvh(cfg["scene"].vertices, cfg["scene"].indices)
    info = ctx.bvh_info(bvh)
    assert info["triangles"] == 9_998_802 and 2 * info["max_depth"] + 1 <= 48, info
    r.ray_log_begin(3_000_000, 1)
    r.render(3)
    log = r.ray_log_end(3_000_000)
    n = len(log["rays"])
    assert n >= 2_000_000, n
    ctx.counters_enable(True)
    ctx.counters_read(True)
    try:
        prim, t, _ = ctx.trace_closest(bvh, log["rays"], log["last_prim"])
        _assert_hits_equal(prim, t, log, "per-thread traversal")
        for flavour in (0, 2, 6):
            ctx.set_trace_flavour(flavour)
            qprim, qt, _ = ctx.trace_queue(bvh, log["rays"], None, log["last_prim"])
            _assert_hits_equal(qprim, qt, log, f"ray-queue kernel, flavour {flavour}")
            assert np.array_equal(qprim, prim) and np.array_equal(qt, t)
        rng = np.random.default_rng(11)
        m = 500_000
        rays, last = log["rays"][:m], log["last_prim"][:m]
        t_max = np.where(log["t"][:m] > 0, log["t"][:m] * rng.uniform(0.5, 1.5, m).astype(np.float32), rng.uniform(0.1, 50.0, m).astype(np.float32)).astype(np.float32)
        occ_ref = r.trace_shadow(rays, t_max, last)
        assert np.array_equal(ctx.trace_shadow(bvh, rays, t_max, last), occ_ref)
        assert np.array_equal(ctx.trace_queue(bvh, rays, t_max, last), occ_ref)
        c = ctx.counters_read(True)
        assert c["stack_overflows"] == 0, c
        # the same rays on the tree the DEVICE builder makes of the scene: hit ids are a property of the triangles, not
        # of the tree, so they must again equal the reference's
        for algorithm in ("ploc", "lbvh"):
            bvh_d = ctx.build_bvh(cfg["scene"].vertices, cfg["scene"].indices, device=algorithm)
            try:
                info_d = ctx.bvh_info(bvh_d)
                print("host build:", info, f"\ndevice build ({algorithm}):", info_d)
                assert info_d["built_on_device"] and info_d["triangles"] == 9_998_802 and 2 * info_d["max_depth"] + 1 <= 48, info_d
                assert info_d["device_build_seconds"] < 1.0, info_d
                assert info_d["sah_cost"] < 1.15 * info["sah_cost"], (algorithm, info_d["sah_cost"], info["sah_cost"])
                dprim, dt, _ = ctx.trace_queue(bvh_d, log["rays"], None, log["last_prim"])
                _assert_hits_equal(dprim, dt, log, f"ray-queue kernel on the device-built tree ({algorithm})")
                assert np.array_equal(ctx.trace_queue(bvh_d, rays, t_max, last), occ_ref)
                assert ctx.counters_read(True)["stack_overflows"] == 0
            finally:
                ctx.destroy_bvh(bvh_d)
    finally:
        ctx.set_trace_flavour(-1)
        ctx.counters_enable(False)
        ctx.destroy_bvh(bvh)


@pytest.mark.parametrize("algorithm", ["ploc", "lbvh"])
@pytest.mark.parametrize("scene_name", ["cornell", "synthetic_small", "tiny"])
def test_device_built_bvh_gives_the_same_hits(pt, wl, ctx, scene_name, algorithm):
    """b200pt_bvh_build with B200PT_BVH_BUILD_DEVICE (Morton order, PLOC clustering or radix tree, SAH-optimal collapse on the GPU) against the host builder's tree:
    closest hits (id, t, uv bits; exact-distance ties excluded) and occlusion flags of random + axis-parallel rays, per-thread traversal and
    ray-queue kernel; edge cases: one and two triangles."""
    rng = np.random.default_rng(17)
    if scene_name == "tiny":
        cases = [(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32), np.array([0, 1, 2], np.int32)),
                 (np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [0, 1, 1]], np.float32), np.array([0, 1, 2, 3, 4, 5], np.int32))]
    else:
        scene = getattr(wl, scene_name)(64, 48)["scene"]
        cases = [(scene.vertices, scene.indices)]
    for vertices, indices in cases:
        host, dev = ctx.build_bvh(vertices, indices), ctx.build_bvh(vertices, indices, device=algorithm)
        try:
            ih, idv = ctx.bvh_info(host), ctx.bvh_info(dev)
            assert idv["built_on_device"] and not ih["built_on_device"] and idv["triangles"] == ih["triangles"] == len(indices) // 3
            lo, hi = vertices.min(axis=0), vertices.max(axis=0)
            n = 200_000
            o = rng.uniform(lo - 0.3 * (hi - lo) - 0.1, hi + 0.3 * (hi - lo) + 0.1, (n, 3)).astype(np.float32)
            d = rng.normal(size=(n, 3)).astype(np.float32)
            d[::11, 1] = 0.0
            d[::13, 2] = -0.0
            d /= np.linalg.norm(d, axis=1, keepdims=True)
            rays = np.ascontiguousarray(np.concatenate([o, d], axis=1), np.float32)
            last = rng.integers(-1, len(indices) // 3, n).astype(np.int32)
            t_max = rng.uniform(0.05, float(np.linalg.norm(hi - lo)) + 1.0, n).astype(np.float32)
            ph, th, uh = ctx.trace_closest(host, rays, last)
            for fn in (ctx.trace_closest, lambda b, r, l: ctx.trace_queue(b, r, None, l)):
                pd, td, ud = fn(dev, rays, last)
                ties = (pd != ph) & (td == th)
                assert ((pd == ph) | ties).all(), f"{int(((pd != ph) & ~ties).sum())} hit ids differ between the device-built and the host-built tree"
                same = pd == ph
                assert np.array_equal(td[same], th[same]) and np.array_equal(ud[same], uh[same])
            assert (ph >= 0).any()
            assert np.array_equal(ctx.trace_shadow(dev, rays, t_max, last), ctx.trace_shadow(host, rays, t_max, last))
            assert np.array_equal(ctx.trace_queue(dev, rays, t_max, last), ctx.trace_shadow(host, rays, t_max, last))
        finally:
            ctx.destroy_bvh(host)
            ctx.destroy_bvh(dev)


def test_c2_size_ray_batches(pt, wl, ctx, ref):
    """SURVEY 8(d) C2 check (i) at its stated size: the 2 073 600 primary rays of the 1920x1080 Cornell frame and >= 1e7 bounce /
    shadow rays the reference's CPU render traces for it, hit ids bit-compared (exact-distance ties excluded)."""
    cfg = wl.cornell(1920, 1080)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    cap = 16_000_000
    r.ray_log_begin(cap, 1)
    r.render(3)
    log = r.ray_log_end(cap)
    n = len(log["rays"])
    assert n >= 2_073_600 + 10_000_000, n
    assert int((log["last_prim"] == -1).sum()) >= 2_073_600          # camera rays leave from no primitive
    ctx.counters_enable(True)
    ctx.counters_read(True)
    try:
        prim, t, _ = ctx.trace_queue(g.bvh, log["rays"], None, log["last_prim"])
        _assert_hits_equal(prim, t, log, "ray-queue kernel")
        assert ctx.counters_read(True)["stack_overflows"] == 0
    finally:
        ctx.counters_enable(False)
    g.close()


def test_frame_parity_per_sample(pt, wl, ctx, ref):
    """Matched RNG streams (CPU seed schedule): G-buffer identical, radiance within the north-star tolerance."""
    cfg = wl.cornell(256, 192)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    spp = 8
    for s in range(spp):
        r.render(1)
        g.render_frame()
        if s == 0:
            g.synchronize()
            for name in ("first_hit_prim_index", "camera_ray_hit", "shading_normals", "geometric_normals", "first_hits", "view_directions"):
                a, b = r.buffer(name), g.download(name)
                if name in ("shading_normals", "geometric_normals", "first_hits"):
                    hit = r.buffer("camera_ray_hit") == 1
                    a, b = a[hit], b[hit]
                assert np.array_equal(a, b), name
            hit = r.buffer("camera_ray_hit") == 1
            assert np.array_equal(r.buffer("materials")[hit][:, 4:244], g.download("materials")[hit][:, 4:244])
            assert np.array_equal(r.buffer("ray_volume_states")[hit][:, :28], g.download("ray_volume_states")[hit][:, :28])
    g.synchronize()
    a, b = r.pixels() / spp, g.download("pixels") / spp
    err = _rel_err(a, b)
    assert err < 1e-3, err          # tolerance stated by BASELINE.json north_star
    assert abs(a.mean() - b.mean()) / a.mean() < 1e-4
    g.close()


@pytest.mark.parametrize("ggx", [0, 1])
def test_envmap_scene_parity(pt, wl, ctx, ref, ggx):
    """Config C3 cut down to parity size: procedural terrain + spheres (64 random principled materials: metals, coats, sheen,
    anisotropy, glass), no emissives, HDR envmap sampled through the alias table with BSDF MIS, both GGX samplers.
    Matched RNG streams: hit ids identical, radiance within the north-star tolerance; phased == megakernel bit for bit."""
    cfg = wl.synthetic_small(160, 96, ggx=ggx)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    spp = 4
    r.render(spp)
    frames = {}
    try:
        for mode in ("megakernel", "phased"):
            ctx.set_path_mode(mode)
            g.reset()
            g.resize(160, 96)
            wl.apply_settings(g.render_settings, cfg["settings"])
            for _ in range(spp):
                g.render_frame()
            g.synchronize()
            frames[mode] = g.download("pixels").copy()
    finally:
        ctx.set_path_mode("phased")
    assert np.array_equal(r.buffer("first_hit_prim_index"), g.download("first_hit_prim_index"))
    assert np.array_equal(frames["megakernel"], frames["phased"])
    a, b = r.pixels() / spp, frames["phased"] / spp
    # the sun (5e4) makes single fireflies dominate a plain mean: compare where the oracle is finite and not a firefly too
    err = _rel_err(a, b)
    assert err < 1e-3, err
    assert abs(a.mean() - b.mean()) / a.mean() < 1e-3
    g.close()


def test_tile_sharding_is_bit_identical_and_deterministic(pt, wl, ctx):
    """Size-independent properties at the full 1080p size: a render split into row tiles (the multi-GPU shard shape)
    equals the unsharded render bit for bit, and re-rendering with the same seeds reproduces the same bits."""
    cfg = wl.cornell(1920, 1080)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    g.render_frame()
    g.render_frame()
    g.synchronize()
    whole = g.download("pixels").copy()
    g.reset()
    g.resize(1920, 1080)
    wl.apply_settings(g.render_settings, cfg["settings"])
    g.render_frame()
    g.render_frame()
    g.synchronize()
    assert np.array_equal(whole, g.download("pixels"))
    # 4 row tiles (multiples of 8 rows), each rendered by separate launches
    g.reset()
    g.resize(1920, 1080)
    wl.apply_settings(g.render_settings, cfg["settings"])
    S = pt.structs
    import ctypes as C
    lib = ctx.lib
    rd = g.render_data
    res = S.Int2(1920, 1080)
    bounds = [0, 272, 544, 816, 1080]
    g.render_frame()     # sample 0 through the host path ...
    g.synchronize()
    first = g.download("pixels").copy()
    g.reset()
    g.resize(1920, 1080)
    wl.apply_settings(g.render_settings, cfg["settings"])
    # ... and the same sample through per-tile launches of the C ABI with the host's render data (seed 42, sample 0)
    host = g.host
    host.b200pt_renderer_set_tile(g.handle, C.byref(S.Tile(0, 0, 0, 0)))   # empty tile: update() only, no pixels touched
    g.render_frame()
    g.synchronize()
    rd = g.render_data
    rd.render_settings.sample_number = 0
    rd.render_settings.need_to_reset = 1
    rd.random_seed = 42
    for y0, y1 in zip(bounds[:-1], bounds[1:]):
        tile = S.Tile(0, y0, 1920, y1)
        pt._capi.check(lib.b200pt_launch_CameraRays(ctx.handle, C.byref(rd), res, C.byref(tile), None), "CameraRays")
        pt._capi.check(lib.b200pt_launch_FullPathTracer(ctx.handle, C.byref(rd), res, C.byref(tile), None), "FullPathTracer")
    ctx.sync()
    assert np.array_equal(first, g.download("pixels"))
    g.close()


@pytest.mark.parametrize("variant", [dict(lss="mis", ess="none"), dict(lss="ris", ess="alias_table")])
def test_path_modes_are_bit_identical(pt, wl, ctx, variant):
    """megakernel / fused / split / phased wavefront shapes of FullPathTracer run the same per-path arithmetic in the same
    RNG order: the accumulated frames must be equal bit for bit (so parity with the oracle carries over to every shape)."""
    cfg = wl.cornell(320, 200)
    cfg["options"] = dict(ggx=0, later=3, **variant)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    frames = {}
    try:
        for mode in ("megakernel", "fused", "split", "phased"):
            ctx.set_path_mode(mode)
            g.reset()
            g.resize(320, 200)
            wl.apply_settings(g.render_settings, cfg["settings"])
            for _ in range(3):
                g.render_frame()
            g.synchronize()
            frames[mode] = g.download("pixels").copy()
    finally:
        ctx.set_path_mode("phased")
        g.close()
    for mode in ("fused", "split", "phased"):
        assert np.array_equal(frames["megakernel"], frames[mode]), mode


@pytest.mark.parametrize("lss,ess", [("ris", "none"), ("ris", "alias_table"), ("restir_di", "none"), ("restir_di", "alias_table")])
def test_ris_phases_are_bit_identical(pt, wl, ctx, lss, ess):
    """Phased RIS (R1 candidates -> closest trace -> R2 select -> [envmap phases] -> any-hit trace -> resolve; under ReSTIR DI from bounce 1
    on) against the megakernel and the split shape on a many-light scene (with a procedural envmap for the alias-table permutations, so that
    emissive triangles AND the envmap are sampled): the per-thread pieces run in the same order on the same generator states, same bits."""
    procedural = importlib.import_module("hiprt-path-tracer_b200.procedural")
    W, H = 160, 112
    cfg = wl.many_lights(W, H, boxes=100, lights=1500)
    cfg["options"] = dict(lss=lss, ess=ess, ggx=0, later=3)
    if ess == "alias_table":
        cfg["envmap"] = {"texels": procedural.procedural_envmap(256, 128), "alias_probas": None, "alias_alias": None, "total_sum": 0.0}
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    frames = {}
    try:
        for mode in ("megakernel", "split", "phased"):
            ctx.set_path_mode(mode)
            g.reset()
            g.resize(W, H)
            wl.apply_settings(g.render_settings, cfg["settings"])
            for _ in range(3):
                g.render_frame()
            g.synchronize()
            frames[mode] = g.download("pixels").copy()
    finally:
        ctx.set_path_mode("phased")
        g.close()
    assert np.isfinite(frames["phased"]).all() and frames["phased"].sum() > 0
    for mode in ("split", "phased"):
        assert np.array_equal(frames["megakernel"], frames[mode]), mode


@pytest.mark.parametrize("ess,neighbors,boost,fused,passes", [("none", 2, True, True, 1), ("alias_table", 2, True, True, 1), ("none", 5, False, True, 1), ("none", 1, False, True, 1),
                                                             ("none", 2, True, True, 2), ("none", 2, True, False, 2), ("alias_table", 3, False, False, 1)])
def test_restir_phases_are_bit_identical(pt, wl, ctx, ess, neighbors, boost, fused, passes):
    """Phased ReSTIR DI passes (InitialCandidates: collect -> closest-hit trace -> resolve -> any-hit trace -> visibility; fused
    SpatiotemporalReuse: collect -> any-hit trace -> resolve, rays through the ray-queue kernel) against the single kernels that trace per
    thread: every frame's initial-candidates and output reservoirs, and the accumulated image, are the same bits over 5 frames (temporal
    history, disocclusion boost, neighbours with and without the visibility term, envmap candidates with the alias-table permutation). The
    stand-alone TemporalReuse / SpatialReuse passes (fused = False, and the spatial passes after the fused one) are phased the same way."""
    procedural = importlib.import_module("hiprt-path-tracer_b200.procedural")
    W, H = 160, 112
    cfg = wl.many_lights(W, H, boxes=100, lights=1500, fused=fused, spatial_passes=passes)
    cfg["options"] = dict(lss="restir_di", ess=ess, ggx=0, later=3)
    if ess == "alias_table":
        cfg["envmap"] = {"texels": procedural.procedural_envmap(256, 128), "alias_probas": None, "alias_alias": None, "total_sum": 0.0}
    cfg["settings"] = dict(cfg["settings"], **{"restir_di_settings.spatial_pass.reuse_neighbor_count": neighbors,
                                               "restir_di_settings.spatial_pass.do_disocclusion_reuse_boost": int(boost),
                                               "restir_di_settings.spatial_pass.neighbor_visibility_count": 3 if neighbors == 5 else 5})
    results = {}
    try:
        for phased in (False, True):
            ctx.set_restir_phases(phased)
            g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="gpu")
            per_frame = []
            for _ in range(5):
                g.render_frame()
                g.synchronize()
                per_frame.append((g.download("restir_initial_candidates").copy(), g.download("restir_output_reservoirs").copy()))
            results[phased] = (per_frame, g.download("pixels").copy())
            g.close()
    finally:
        ctx.set_restir_phases(True)
    assert np.isfinite(results[True][1]).all() and results[True][1].sum() > 0
    for frame, (single, phased) in enumerate(zip(results[False][0], results[True][0])):
        assert np.array_equal(single[0], phased[0]), f"initial candidates differ in frame {frame}"
        assert np.array_equal(single[1], phased[1]), f"output reservoirs differ in frame {frame}"
    assert np.array_equal(results[False][1], results[True][1])
    out = _reservoirs(results[True][0][-1][1])
    assert (out["M"] > 1).mean() > 0.5          # the temporal history was reused


def test_edge_cases_empty_scene_and_misses(pt, ctx):
    """Empty scene (0 triangles): every query misses; rays pointing away from the geometry miss; zero-length batches work."""
    bvh = ctx.build_bvh(np.zeros((0, 3), np.float32), np.zeros((0,), np.int32))
    rays = np.array([[0, 0, 0, 0, 0, 1], [1, 2, 3, 1, 0, 0]], np.float32)
    prim, t, _ = ctx.trace_closest(bvh, rays)
    assert (prim == -1).all() and (t == -1.0).all()
    assert (ctx.trace_shadow(bvh, rays, np.array([10.0, 10.0], np.float32)) == 0).all()
    ctx.destroy_bvh(bvh)
    verts = np.array([[0, 0, 5], [1, 0, 5], [0, 1, 5]], np.float32)
    bvh = ctx.build_bvh(verts, np.array([0, 1, 2], np.int32))
    rays = np.array([[0.2, 0.2, 0, 0, 0, 1], [0.2, 0.2, 0, 0, 0, -1], [5, 5, 0, 0, 0, 1], [0.2, 0.2, 0, 0, 0, 1]], np.float32)
    prim, t, uv = ctx.trace_closest(bvh, rays, np.array([-1, -1, -1, 0], np.int32))
    assert prim.tolist() == [0, -1, -1, -1]          # hit, behind, beside, filtered as the last-hit primitive
    assert t[0] == 5.0 and np.allclose(uv[0], [0.2, 0.2])
    occ = ctx.trace_shadow(bvh, rays[:1].repeat(3, 0), np.array([4.0, 5.0, 5.001], np.float32))
    assert occ.tolist() == [0, 0, 1]                 # occluded iff t < t_max - 1e-4
    p0, t0, _ = ctx.trace_closest(bvh, np.zeros((0, 6), np.float32))
    assert len(p0) == 0
    ctx.destroy_bvh(bvh)


def test_uncompiled_variant_is_refused(pt, ctx):
    g = pt.GPURenderer(ctx)
    with pytest.raises(pt.B200ptError, match="not compiled"):
        g.set_kernel_options(lss="bsdf", ess="binary_search", ggx=1)
    g.close()


RESERVOIR = np.dtype([("M", "<i4"), ("weight_sum", "<f4"), ("UCW", "<f4"), ("tri", "<i4"), ("point", "<f4", 3),
                      ("target", "<f4"), ("flags", "u1"), ("pad", "u1", 3)])
PRESAMPLED = np.dtype([("tri", "<i4"), ("point", "<f4", 3), ("normal", "<f4", 3), ("radiance", "<f4", 3), ("pdf", "<f4"),
                       ("flags", "u1"), ("pad", "u1", 3)])


def _reservoirs(raw):
    return np.ascontiguousarray(raw).view(RESERVOIR).reshape(-1)


def test_restir_di_parity(pt, wl, ctx, ref):
    """Config C4 cut down to parity size: many-light room, LSS_RESTIR_DI with presampled lights, 4+1 initial candidates with
    visibility reuse, fused spatiotemporal reuse (pairwise-defensive MIS, jacobian, heuristics), RIS on later bounces.
    Matched RNG streams over 4 frames so that the temporal ping-pong of the reservoir buffers is exercised.
    Integer work (which light each reservoir holds, M) is compared exactly; a reservoir may legitimately differ where a
    libm sinf/cosf ulp difference moved a disk-sampled neighbour by a pixel, so a small mismatch budget is allowed."""
    cfg = wl.many_lights(128, 96, boxes=150, lights=3000)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    frames = 4
    n = 128 * 96
    for f in range(frames):
        r.render(1)
        g.render_frame()
        g.synchronize()
        # presampled lights: no trigonometry on this path -> byte-exact
        pa = np.ascontiguousarray(r.buffer("presampled_lights")).view(PRESAMPLED).reshape(-1)
        pb = np.ascontiguousarray(g.download("presampled_lights")).view(PRESAMPLED).reshape(-1)
        for field in ("tri", "point", "normal", "radiance", "pdf", "flags"):
            assert np.array_equal(pa[field], pb[field]), (f, field)
        ia, ib = _reservoirs(r.buffer("restir_initial_candidates")), _reservoirs(g.download("restir_initial_candidates"))
        assert np.array_equal(ia["M"], ib["M"]), f
        same = (ia["tri"] == ib["tri"]) & (ia["flags"] == ib["flags"])
        assert same.mean() > 0.999, (f, same.mean())
        assert np.allclose(ia["UCW"][same], ib["UCW"][same], rtol=2e-3, atol=1e-12), f
        oa, ob = _reservoirs(r.buffer("restir_output_reservoirs")), _reservoirs(g.download("restir_output_reservoirs"))
        same = (oa["tri"] == ob["tri"]) & (oa["M"] == ob["M"])
        assert same.mean() > 0.995, (f, same.mean())
        if f < 2:
            assert same.all(), f        # nothing has had a chance to cascade yet: every reservoir holds the oracle's light
        ok = same & np.isfinite(oa["UCW"])
        close = np.isclose(oa["UCW"][ok], ob["UCW"][ok], rtol=1e-4, atol=1e-12)
        assert close.mean() > 0.995, (f, close.mean())
    assert ia.shape[0] == n
    a, b = r.pixels() / frames, g.download("pixels") / frames
    err = _rel_err(a, b)
    assert err < 1e-3, err
    assert abs(a.mean() - b.mean()) / a.mean() < 1e-3
    g.close()


@pytest.mark.parametrize("lss,ess,ggx,oracle_variant", [
    ("mis", "none", 1, "mis_caps"), ("ris", "none", 1, "ris_caps"), ("ris", "alias_table", 1, "ris_caps"),
    ("restir_di", "none", 1, "restir_caps"), ("restir_di", "alias_table", 0, "restir_alias"), ("restir_di", "alias_table", 1, "restir_alias_caps")])
def test_kernel_permutation_matrix_parity(pt, wl, ctx, ref, lss, ess, ggx, oracle_variant):
    """{MIS, RIS, ReSTIR DI} x {no envmap sampling, alias table} x {VNDF, spherical caps}. The permutations the configs do
    not use, each against the reference's CPU kernels compiled with the same KernelOptions (oracle/build_ref.py), matched RNG streams,
    3 frames, mean relative radiance error <= 1e-3 (north-star tolerance) and G-buffer hit ids identical."""
    if not ref.available(oracle_variant):
        pytest.skip(f"oracle variant {oracle_variant} is not built")
    procedural = importlib.import_module("hiprt-path-tracer_b200.procedural")
    if lss == "restir_di":
        cfg = wl.many_lights(112, 80, boxes=120, lights=2000)
    elif ess == "alias_table":
        cfg = wl.synthetic_small(112, 80, ggx=ggx)
    else:
        cfg = wl.cornell(112, 80)
    if ess == "alias_table" and cfg["envmap"] is None:
        cfg["envmap"] = {"texels": procedural.procedural_envmap(256, 128), "alias_probas": None, "alias_alias": None, "total_sum": 0.0}
    cfg["options"] = dict(lss=lss, ess=ess, ggx=ggx, later=3)
    cfg["oracle_variant"] = oracle_variant
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    frames = 3
    for _ in range(frames):
        r.render(1)
        g.render_frame()
    g.synchronize()
    assert np.array_equal(g.download("first_hit_prim_index"), r.buffer("first_hit_prim_index"))
    a, b = r.pixels() / frames, g.download("pixels") / frames
    assert np.isfinite(b).all() and a.mean() > 0
    err = _rel_err(a, b)
    assert err < 1e-3, err
    assert abs(a.mean() - b.mean()) / a.mean() < 1e-3
    g.close()


@pytest.mark.parametrize("temporal,spatial,passes", [(True, True, 1), (True, True, 2), (False, True, 2), (True, False, 1)])
def test_restir_di_separate_reuse_passes_parity(pt, wl, ctx, ref, temporal, spatial, passes):
    """the stand-alone ReSTIR_DI_TemporalReuse / ReSTIR_DI_SpatialReuse kernels (do_fused_spatiotemporal off) and
    the host ping-pong of their reservoir buffers (temporal pass in place over the initial candidates, spatial passes between the
    two spatial buffers), against the reference's CPU kernels driven the way CPURenderer::ReSTIR_DI drives them
    (CPURenderer.cpp:413-429). Same comparison rules as the fused test: integers exact, a small budget for reservoirs where a
    sinf/cosf ulp moved a disk-sampled neighbour by a pixel."""
    cfg = wl.many_lights(128, 96, boxes=150, lights=3000, fused=False, temporal=temporal, spatial=spatial, spatial_passes=passes)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    frames = 4
    import ctypes as C
    try:
        for f in range(frames):
            r.render(1)
            g.render_frame()
            g.synchronize()
            raw = np.ascontiguousarray(r.buffer("restir_output_reservoirs"))
            oa, ob = _reservoirs(raw), _reservoirs(g.download("restir_output_reservoirs"))
            same = (oa["tri"] == ob["tri"]) & (oa["M"] == ob["M"])
            assert same.mean() > 0.998, (f, same.mean())
            ok = same & np.isfinite(oa["UCW"])
            close = np.isclose(oa["UCW"][ok], ob["UCW"][ok], rtol=1e-4, atol=1e-12)
            assert close.mean() > 0.995, (f, close.mean())
            if f > 0 or spatial:
                assert (ob["M"] > 1).mean() > 0.3, (f, (ob["M"] > 1).mean())      # the passes did combine reservoirs
            # A reservoir that differs (a sinf/cosf ulp moved a disk-sampled neighbour by a pixel) spreads over reuse_radius per
            # spatial pass and per frame; the oracle's output becomes the next frame's temporal input on both sides so that
            # every frame is checked on its own
            pt._capi.check(ctx.lib.b200pt_memcpy_h2d(ctx.handle, C.c_void_p(g.device_pointer("restir_output_reservoirs")), raw.ctypes.data_as(C.c_void_p), C.c_size_t(raw.nbytes), None), "h2d")
            ctx.sync()
        a, b = r.pixels() / frames, g.download("pixels") / frames
        err = _rel_err(a, b)
        assert err < 1e-3, err
        assert abs(a.mean() - b.mean()) / a.mean() < 1e-3
    finally:
        g.close()


@pytest.mark.parametrize("temporal,spatial,passes", [(True, True, 1), (True, True, 2)])
def test_restir_di_separate_reuse_passes_own_history(pt, wl, ctx, ref, temporal, spatial, passes):
    """The same non-fused configuration WITHOUT handing the oracle's reservoirs to the GPU between frames: the temporal pass of frame k reads what
    the GPU's own passes left in frame k - 1 (the host ping-pong of ReSTIRDIRenderPass.cpp:332-425 carried over several frames). A reservoir that
    differs (a sinf / cosf ulp moves a disk-sampled neighbour by a pixel) now spreads by the reuse radius per pass and frame, so the share of equal
    reservoirs is bounded per frame instead of pinned, as in the fused test; the accumulated image stays within the north-star tolerance."""
    cfg = wl.many_lights(128, 96, boxes=150, lights=3000, fused=False, temporal=temporal, spatial=spatial, spatial_passes=passes)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    frames = 5
    shares, tight = [], []
    try:
        for f in range(frames):
            r.render(1)
            g.render_frame()
            g.synchronize()
            oa, ob = _reservoirs(np.ascontiguousarray(r.buffer("restir_output_reservoirs"))), _reservoirs(g.download("restir_output_reservoirs"))
            same = (oa["tri"] == ob["tri"]) & (oa["M"] == ob["M"])
            shares.append(float(same.mean()))
            assert same.mean() > (0.998 if f == 0 else 0.97), (f, shares)
            ok = same & np.isfinite(oa["UCW"])
            # the weights carry the float sums of every earlier frame's neighbours: one reservoir that differs changes the weight sums (not the
            # selected light) of everything within the reuse radius, pass after pass -- on B200 the share of weights within 1e-4 of the reference's
            # goes 1.0, 0.9995, 0.994, 0.959, 0.875 over five frames of the two-pass shape while the lights stay equal on 99.95 % of the pixels.
            # Bounded loosely per frame, pinned on the first; the accumulated image below is the real check
            tight.append(float(np.isclose(oa["UCW"][ok], ob["UCW"][ok], rtol=1e-4, atol=1e-12).mean()))
            loose = float(np.isclose(oa["UCW"][ok], ob["UCW"][ok], rtol=5e-2, atol=1e-12).mean())
            assert (f > 0 or tight[-1] > 0.995) and loose > 0.98, (f, shares, tight, loose)
        assert (ob["M"] > 4).mean() > 0.3, (ob["M"] > 4).mean()          # history did accumulate over the frames
        print("share of reservoirs equal to the reference's, per frame, own history:", shares, "weights within 1e-4:", tight)
        a, b = r.pixels() / frames, g.download("pixels") / frames
        err = _rel_err(a, b)
        assert err < 1e-3, err
        assert abs(a.mean() - b.mean()) / a.mean() < 1e-3
    finally:
        g.close()


def test_restir_di_extra_spatial_pass_after_fused(pt, wl, ctx):
    """GPU schedule of the reference (ReSTIRDIRenderPass.cpp:476-503): number_of_passes = 2 runs one SpatialReuse pass after the
    fused spatiotemporal pass. The CPU renderer never does (CPURenderer.cpp:405-411), so there is no oracle run for this shape:
    checked here are the buffer ping-pong (the output of frame k is the temporal input of frame k + 1, M grows to the cap),
    determinism, and that the estimate agrees with the one-pass render of the same scene (both are unbiased: pairwise MIS)."""
    W, H, frames = 128, 96, 24
    means = []
    for passes in (1, 2):
        cfg = wl.many_lights(W, H, boxes=150, lights=3000, spatial_passes=passes)
        runs = []
        for _ in range(2 if passes == 2 else 1):
            g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="gpu")
            for _ in range(frames):
                g.render_frame()
            g.synchronize()
            runs.append((g.download("pixels").copy() / frames, _reservoirs(g.download("restir_output_reservoirs")).copy()))
            g.close()
        if passes == 2:
            assert np.array_equal(runs[0][0], runs[1][0])
            M, cap = runs[0][1]["M"], 25       # ReSTIRDISettings.h m_cap default
            assert M.max() == cap and (M == cap).mean() > 0.5, (M.max(), (M == cap).mean())     # m_cap: the history survived every frame
        means.append(runs[0][0].mean())
    assert abs(means[0] - means[1]) / means[0] < 0.02, means


def test_tonemap_and_headless_png_parity(pt, wl, ctx, ref, tmp_path):
    """the headless output path. b200pt_tonemap against CPURenderer::tonemap + write_image_png's quantisation
    (the reference's own ColorRGB32F exp / pow through the oracle harness) on the same accumulated frame, then the PNG written
    from it decodes to the same bytes. expf / powf differ by an ulp between glibc and CUDA: floats within 2e-6, and an 8-bit
    value may land on the other side of a truncation boundary in a handful of pixels (never by more than 1)."""
    io = importlib.import_module("hiprt-path-tracer_b200.image_io")
    cfg = wl.cornell(160, 96)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    try:
        spp = 3
        r.render(spp)
        for _ in range(spp):
            g.render_frame()
        g.synchronize()
        # the same accumulated frame on both sides (single pixels of the two renders may differ by more than an ulp; the frame
        # parity tests bound that): the oracle's accumulation buffer replaces the GPU's
        import ctypes as C
        acc = np.ascontiguousarray(r.pixels())
        assert _rel_err(acc, g.download("pixels")) < 1e-3
        pt._capi.check(ctx.lib.b200pt_memcpy_h2d(ctx.handle, C.c_void_p(g.device_pointer("pixels")), acc.ctypes.data_as(C.c_void_p), C.c_size_t(acc.nbytes), None), "h2d")
        ctx.sync()
        want_f, want_u8 = r.tonemap(2.2, 1.0)
        got_f, got_u8 = g.tonemap(2.2, 1.0)
        # 1 - exp(-c) cancels for near-black pixels: an ulp of expf (6e-8 near 1) is the whole value there and pow(., 1 / 2.2)
        # magnifies it (0 vs 1.2e-7 -> 0 vs 7e-4). Compare before the gamma curve, where the error is the ulp of expf itself.
        assert np.abs(want_f.astype(np.float64) ** 2.2 - got_f.astype(np.float64) ** 2.2).max() < 2.5e-7
        assert np.abs(want_f - got_f).max() < 1e-3 and np.abs(want_f - got_f).mean() < 1e-6
        d = np.abs(want_u8.astype(np.int32) - got_u8.astype(np.int32))
        assert d.max() <= 1 and (d != 0).mean() < 2e-3, (d.max(), (d != 0).mean())
        assert got_u8.max() > 200 and got_u8.min() < 30       # a real image, not a constant
        flipped_f, flipped_u8 = g.tonemap(2.2, 1.0, flip_y=True)
        assert np.array_equal(flipped_u8, got_u8[::-1]) and np.array_equal(flipped_f, got_f[::-1])
        path = os.path.join(tmp_path, "cornell.png")
        io.write_png(path, got_u8)
        assert np.array_equal(io.read_png(path), got_u8)
    finally:
        g.close()


def test_c5_bench_workload_all_features_at_once(pt, wl, ctx, ref):
    """bench.py --workload adaptive: nested dielectrics + thin film + dispersion + adaptive sampling together. Each feature is
    pinned tightly on its own above; combined, the reference itself produces a few NaN samples per frame on the coated dispersive
    glass (it prints 'NaN at ...' and drops them) at pixels that depend on the host's libm, and adaptive sampling then freezes
    those pixels at different samples -- so this test checks what does not depend on that: the two GPU path shapes agree bit for
    bit, the frame is finite, and it agrees with the reference statistically."""
    W, H, spp = 128, 96, 12
    cfg = wl.adaptive(W, H)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    try:
        r.render(spp)
        frames = {}
        for mode in ("megakernel", "phased"):
            ctx.set_path_mode(mode)
            g.reset()
            g.resize(W, H)
            wl.apply_settings(g.render_settings, cfg["settings"])
            for _ in range(spp):
                g.render_frame()
            g.synchronize()
            frames[mode] = g.download("pixels").copy()
        assert np.array_equal(frames["megakernel"], frames["phased"])
        a, b = r.pixels(), frames["phased"]
        assert np.isfinite(b).all()
        conv = g.download("pixel_converged_sample_count")
        assert (conv != -1).mean() > 0.05                      # adaptive sampling did stop pixels
        # Undefined in the reference: thin film on glass lets principled_glass_sample attempt a refraction
        # under total internal reflection and return its UNINITIALISED direction (Principled.h:393-484, Sampling.h:153-166); the
        # oracle build returns a stale direction there, this implementation ends the path: ~5 % of the pixels, frame mean -7 %.
        assert abs(a.mean() - b.mean()) / a.mean() < 0.15
        assert np.median(np.abs(a - b) / np.maximum(a, 1e-3)) < 1e-3
    finally:
        ctx.set_path_mode("phased")
        g.close()


def test_c5_as_specified_defined_part_parity(pt, wl, ctx, ref):
    """SURVEY 8(d) C5 as specified -- nested-dielectrics-complex UNION multi-dispersion (12 090 triangles), procedural envmap through the
    alias table with BSDF MIS, dielectric priorities 1..3 + absorption colours, dispersion, thin film, adaptive sampling at the reference's
    defaults (min 64 samples, threshold 0.3) -- with the one combination left out whose result is undefined in the reference (thin film ON
    glass under total internal reflection returns an uninitialised direction, Principled.h:437, Sampling.h:153-166):
    the coating stays on the non-transmissive materials. Matched streams, 70 samples (adaptive sampling acts from sample 65): radiance
    per pixel sample within 1e-3 mean relative error, converged-pixel counters equal on > 99 % of the pixels."""
    W, H, spp = 80, 56, 70
    cfg = wl.adaptive_full(W, H, envmap_size=(256, 128), thin_film_on_glass=False)
    assert cfg["scene"].triangle_count == 12090
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    try:
        r.render(spp)
        for _ in range(spp):
            g.render_frame()
        g.synchronize()
        ca, cb = r.buffer("pixel_sample_count").astype(np.float64), g.download("pixel_sample_count").astype(np.float64)
        assert (ca == cb).mean() > 0.99, (ca == cb).mean()
        conv_a, conv_b = r.buffer("pixel_converged_sample_count"), g.download("pixel_converged_sample_count")
        assert (conv_a == conv_b).mean() > 0.99 and (conv_b != -1).any(), "adaptive sampling must have stopped some pixels, the same ones"
        same = ca == cb
        a = r.pixels() / np.maximum(ca, 1.0)[..., None]
        b = g.download("pixels") / np.maximum(cb, 1.0)[..., None]
        assert np.isfinite(b).all()
        err = _rel_err(a[same], b[same])
        assert err < 1e-3, err
        assert abs(a.mean() - b.mean()) / a.mean() < 1e-3
    finally:
        g.close()


def _textured_cornell(wl, width, height):
    """Cornell box with 8-bit textures driving the material parameters (get_intersection_material's overrides,
    Material.h:47-96; sample_texture_rgba, Texture.h:30-76; normal mapping): planar texcoords that leave [0, 1] (wrap), an sRGB
    RGBA base colour, a packed roughness-metallic map, single-parameter maps and a normal map on different materials."""
    cfg = wl.cornell(width, height)
    scene = cfg["scene"]
    v = scene.vertices
    scene.texcoords = np.stack([v[:, 0] * 0.37 + 0.5 * v[:, 1], v[:, 2] * 0.41 - 0.3 * v[:, 1]], axis=1).astype(np.float32)
    rng = np.random.default_rng(11)
    def tex(h, w, c, lo=0, hi=256):
        return rng.integers(lo, hi, size=(h, w, c), dtype=np.uint8)
    normal_map = tex(8, 8, 3, 96, 160)
    normal_map[..., 2] = 255                                   # mostly +z: a perturbed, not a random, normal
    scene.textures = [tex(16, 32, 4, 40, 256), tex(16, 16, 3), tex(8, 8, 3, 20, 230), normal_map, tex(4, 4, 3)]
    emissive_materials = set(int(scene.material_indices[t]) for t in scene.emissive)
    plain = [i for i in range(len(scene.materials)) if i not in emissive_materials]
    assert len(plain) >= 5
    m = scene.materials
    m[plain[0]].base_color_texture_index = 0
    m[plain[1]].roughness_metallic_texture_index = 1
    m[plain[2]].roughness_texture_index = 2
    m[plain[2]].specular_texture_index = 4
    m[plain[2]].coat_texture_index = 2
    m[plain[3]].normal_map_texture_index = 3
    m[plain[4]].base_color_texture_index = 0
    m[plain[4]].specular_color_texture_index = 4
    m[plain[4]].sheen_texture_index = 2
    m[plain[4]].sheen_color_texture_index = 4
    cfg["name"] += ", textured"
    return cfg


def test_textured_materials_parity(pt, wl, ctx, ref):
    """Row M against the reference CPU kernels with matched RNG streams: the G-buffer materials (every overridden parameter) and
    shading normals (normal map) are compared byte for byte, radiance within the north-star tolerance, and the wavefront (textured
    hits keep their material in the path-state slot, plain hits keep the index) equals the megakernel bit for bit."""
    W, H, spp = 160, 96, 4
    cfg = _textured_cornell(wl, W, H)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    try:
        r.render(spp)
        frames = {}
        for mode in ("megakernel", "phased"):
            ctx.set_path_mode(mode)
            g.reset()
            g.resize(W, H)
            wl.apply_settings(g.render_settings, cfg["settings"])
            for _ in range(spp):
                g.render_frame()
            g.synchronize()
            frames[mode] = g.download("pixels").copy()
        hit = r.buffer("camera_ray_hit") == 1
        assert np.array_equal(r.buffer("first_hit_prim_index"), g.download("first_hit_prim_index"))
        ma, mb = r.buffer("materials")[hit][:, 4:244].copy(), g.download("materials")[hit][:, 4:244].copy()
        # the sRGB base colour goes through powf(texel, 2.2) (Texture.h:72-73): glibc and CUDA differ by one ulp on ~7 % of the
        # textured pixels; every other parameter (linear texel / 255) is byte-equal
        o = pt.structs.SimplifiedRendererMaterial.base_color.offset - 4
        ca, cb = ma[:, o:o + 12].copy().view(np.float32), mb[:, o:o + 12].copy().view(np.float32)
        assert np.allclose(ca, cb, rtol=1.5e-7, atol=0.0)
        ma[:, o:o + 12] = 0
        mb[:, o:o + 12] = 0
        assert np.array_equal(ma, mb)
        ma[:, o:o + 12] = ca.view(np.uint8)
        assert np.array_equal(r.buffer("shading_normals")[hit], g.download("shading_normals")[hit])
        textured_pixels = (ma.view(np.float32) != ma.view(np.float32)[0]).any(axis=1).mean()
        assert textured_pixels > 0.5                               # the textures do vary the materials over the image
        assert np.array_equal(frames["megakernel"], frames["phased"])
        a, b = r.pixels() / spp, frames["phased"] / spp
        err = _rel_err(a, b)
        assert err < 1e-3, err
    finally:
        ctx.set_path_mode("phased")
        g.close()


@pytest.mark.parametrize("workload", ["textured_cornell", "synthetic_small", "many_lights", "nested_dielectrics"])
def test_shading_records_are_bit_identical(pt, wl, ctx, workload, monkeypatch):
    """b200pt_bvh_attach_shading_data: the per-triangle-slot records (three vertex normals + texcoords in 64 bytes, material index and the
    has-normals flag in the spare words of the triangle record) hold the very values a hit otherwise gathers through the index buffer, so
    every buffer of a render -- G-buffer (materials from textures, normal-mapped shading normals), accumulation, reservoirs -- is equal bit
    for bit with the records attached (the host layer's default) and without (B200PT_SHADING_RECORDS=0), in every path shape."""
    W, H, frames = 160, 96, 3
    def make():
        if workload == "textured_cornell":
            return _textured_cornell(wl, W, H)
        if workload == "synthetic_small":
            return wl.synthetic_small(W, H)
        if workload == "many_lights":
            return wl.many_lights(W, H, boxes=150, lights=3000)
        return wl.golden_scene("nested-dielectrics-complex", W, H)
    outputs = {}
    for records in ("1", "0"):
        monkeypatch.setenv("B200PT_SHADING_RECORDS", records)
        cfg = make()
        for mode in ("phased", "megakernel") if workload != "many_lights" else ("phased",):
            ctx.set_path_mode(mode)
            g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="gpu")
            try:
                for _ in range(frames):
                    g.render_frame()
                g.synchronize()
                names = ["pixels", "first_hit_prim_index", "shading_normals", "geometric_normals", "materials", "denoiser_normals", "denoiser_albedo"]
                if workload == "many_lights":
                    names.append("restir_output_reservoirs")
                outputs[(records, mode)] = {n: g.download(n).copy() for n in names}
            finally:
                g.close()
                ctx.set_path_mode("phased")
    for (records, mode), out in outputs.items():
        if records == "0":
            for name, value in out.items():
                assert np.array_equal(value, outputs[("1", mode)][name], equal_nan=(value.dtype.kind == "f")), (workload, mode, name)
    assert outputs[("1", "phased")]["pixels"].any()


@pytest.mark.parametrize("workload", ["cornell", "synthetic_small", "many_lights", "cornell_ris"])
def test_path_record_layouts_are_bit_identical(pt, wl, ctx, workload, monkeypatch):
    """WavefrontState::stride_*: the seven per-path float4 records grouped into ray / hit / radiance records of 32 / 64 / 32 bytes (default) against seven separate
    arrays (B200PT_PATH_RECORDS=0, round 1's layout) and against the layout that also groups the NEE and hit records (=2). Only addresses change: every path shape (phased with sample batches, split, fused) and the
    ray-queue kernel's strided view of origin / direction give the same bits."""
    W, H, frames = 160, 96, 3
    def make():
        if workload == "cornell":
            return wl.cornell(W, H)
        if workload == "cornell_ris":
            cfg = wl.cornell(W, H)
            cfg["options"] = dict(lss="ris", ess="none", ggx=0, later=3)
            return cfg
        if workload == "synthetic_small":
            return wl.synthetic_small(W, H)
        return wl.many_lights(W, H, boxes=150, lights=3000)
    outputs = {}
    try:
        for layout in ("1", "0", "2", "1"):
            monkeypatch.setenv("B200PT_PATH_RECORDS", layout)
            for mode in ("phased", "split", "fused") if workload != "many_lights" else ("phased",):
                ctx.set_path_mode(mode)
                cfg = make()
                cfg["settings"] = dict(cfg["settings"], samples_per_frame=2)
                g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="gpu")
                try:
                    for _ in range(frames):
                        g.render_frame()
                    g.synchronize()
                    out = {n: g.download(n).copy() for n in ("pixels", "first_hit_prim_index", "shading_normals", "denoiser_albedo")}
                finally:
                    g.close()
                key = (layout, mode)
                if key in outputs:       # the second "1" run: switching back works too
                    for n, v in out.items():
                        assert np.array_equal(v, outputs[key][n], equal_nan=(v.dtype.kind == "f")), (workload, mode, n, "second run")
                outputs[key] = out
    finally:
        ctx.set_path_mode("phased")
        monkeypatch.delenv("B200PT_PATH_RECORDS", raising=False)
    for (layout, mode), out in outputs.items():
        if layout != "1":
            for n, v in out.items():
                assert np.array_equal(v, outputs[("1", mode)][n], equal_nan=(v.dtype.kind == "f")), (workload, mode, n, layout)
    assert outputs[("1", "phased")]["pixels"].any()


def test_column_tile_sharding_is_bit_identical(pt, wl, ctx):
    """The strong-scaling shard shape of config C3 (bench.py --sharding columns) on one device: three renderers own three column
    bands of the envmap scene; the union of their accumulation buffers equals the unsharded render bit for bit."""
    mg = importlib.import_module("hiprt-path-tracer_b200.multigpu")
    W, H, frames = 200, 96, 3
    cfg = wl.synthetic_small(W, H)
    whole = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    try:
        for _ in range(frames):
            whole.render_frame()
        whole.synchronize()
        want = whole.download("pixels").reshape(H, W, 3).copy()
    finally:
        whole.close()
    tiles = mg.column_tiles(W, 3)
    assert tiles == [(0, 64), (64, 128), (128, 200)]
    got = np.zeros_like(want)
    for x0, x1 in tiles:
        g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
        try:
            g.set_tile((x0, 0, x1, H))
            for _ in range(frames):
                g.render_frame()
            g.synchronize()
            band = g.download("pixels").reshape(H, W, 3)
            assert not band[:, :x0].any() and not band[:, x1:].any()      # a rank writes only its own columns
            got[:, x0:x1] = band[:, x0:x1]
        finally:
            g.close()
    assert np.array_equal(got, want)


def test_restir_di_tile_sharding_with_halo_is_bit_identical(pt, wl, ctx):
    """The multi-GPU shard shape of config C4 on one device: two renderers own the two halves of the image, each runs CameraRays
    over tile + 20 halo rows, and after every frame only the output reservoirs of the halo rows are copied across (what
    multigpu.exchange_halo does over NCCL). The union of the tiles must equal the unsharded render bit for bit."""
    import ctypes as C
    mg = importlib.import_module("hiprt-path-tracer_b200.multigpu")
    W, H, frames = 128, 96, 4
    cfg = wl.many_lights(W, H, boxes=150, lights=3000)
    whole = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    for _ in range(frames):
        whole.render_frame()
    whole.synchronize()
    want = whole.download("pixels").reshape(H, W, 3).copy()
    want_res = whole.download("restir_output_reservoirs").reshape(H, W, 36).copy()
    whole.close()

    tiles = mg.row_tiles(H, 2)
    halo = mg.restir_halo_rows()
    parts = []
    for y0, y1 in tiles:
        g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
        g.set_tile((0, y0, W, y1), halo=halo)
        parts.append(g)
    row_bytes = W * 36
    for _ in range(frames):
        for g in parts:
            g.render_frame()
        for g in parts:
            g.synchronize()
        for rank, g in enumerate(parts):
            for peer, send, recv in mg.halo_plan(tiles, rank, halo):
                if recv:
                    src = parts[peer].device_pointer("restir_output_reservoirs") + recv[0] * row_bytes
                    dst = g.device_pointer("restir_output_reservoirs") + recv[0] * row_bytes
                    pt._capi.check(ctx.lib.b200pt_memcpy_d2d(ctx.handle, C.c_void_p(dst), C.c_void_p(src), C.c_size_t((recv[1] - recv[0]) * row_bytes), None), "d2d")
        ctx.sync()
    for (y0, y1), g in zip(tiles, parts):
        got = g.download("pixels").reshape(H, W, 3)
        got_res = g.download("restir_output_reservoirs").reshape(H, W, 36)
        assert np.array_equal(got_res[y0:y1], want_res[y0:y1]), (y0, y1)
        assert np.array_equal(got[y0:y1], want[y0:y1]), (y0, y1)
        g.close()


@pytest.mark.parametrize("name,kwargs", [
    ("nested-dielectrics", dict()),
    ("nested-dielectrics-complex", dict(nb_bounces=12)),
    ("multi-dispersion", dict()),
    ("cornell_pbr", dict(thin_film=True, adaptive=True)),
])
def test_c5_feature_scenes_parity(pt, wl, ctx, ref, name, kwargs):
    """The material features config C5 names, on the reference's own test scenes with matched RNG streams: nested dielectrics
    with the priority interior stack, spectral dispersion (one wavelength per path), thin-film interference, per-pixel adaptive
    sampling. G-buffer hit ids and the adaptive-sampling counters are integer work -> exact; radiance within the north-star
    tolerance; the phased wavefront equals the megakernel bit for bit on each of them."""
    W, H, spp = 128, 96, 12 if kwargs.get("adaptive") else 4
    cfg = wl.golden_scene(name, W, H, **kwargs)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    r.render(spp)
    frames = {}
    try:
        for mode in ("megakernel", "phased"):
            ctx.set_path_mode(mode)
            g.reset()
            g.resize(W, H)
            wl.apply_settings(g.render_settings, cfg["settings"])
            for _ in range(spp):
                g.render_frame()
            g.synchronize()
            frames[mode] = g.download("pixels").copy()
    finally:
        ctx.set_path_mode("phased")
    assert np.array_equal(frames["megakernel"], frames["phased"])
    prim_a, prim_b = r.buffer("first_hit_prim_index"), g.download("first_hit_prim_index")
    if kwargs.get("adaptive"):
        conv_a, conv_b = r.buffer("pixel_converged_sample_count"), g.download("pixel_converged_sample_count")
        assert (conv_a != conv_b).mean() < 2e-3                     # a pixel sitting on the noise threshold may flip on an ulp
        assert (conv_a != -1).any()                                 # the test did exercise convergence
        # a converged pixel keeps the G-buffer of the last sample it was active in: hit ids and sample counts are compared
        # where both sides agree on when (or that not) the pixel converged
        same = conv_a == conv_b
        assert np.array_equal(prim_a[same], prim_b[same])
        assert np.array_equal(r.buffer("pixel_sample_count")[same], g.download("pixel_sample_count")[same])
    else:
        assert np.array_equal(prim_a, prim_b)
    a, b = r.pixels() / spp, frames["phased"] / spp
    err = _rel_err(a, b)
    assert err < 1e-3, err
    assert abs(a.mean() - b.mean()) / a.mean() < 1e-3
    g.close()


def test_path_lanes_are_bit_identical(pt, wl, ctx):
    """FullPathTracer as 1, 2 or 3 concurrent band launches on separate streams (the default is 2) gives the same frame."""
    cfg = wl.cornell(1920, 1080)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    frames = {}
    for lanes in (1, 2, 3):
        g.set_path_lanes(lanes)
        assert g.path_lanes == lanes
        g.reset()
        g.resize(1920, 1080)
        wl.apply_settings(g.render_settings, cfg["settings"])
        for _ in range(3):
            g.render_frame()
        g.synchronize()
        frames[lanes] = g.download("pixels").copy()
    assert np.array_equal(frames[1], frames[2])
    assert np.array_equal(frames[1], frames[3])
    g.close()


@pytest.mark.parametrize("workload,schedule", [("cornell", "gpu"), ("cornell", "cpu"), ("synthetic_small", "gpu"), ("cornell_ris", "gpu")])
def test_sample_batches_are_bit_identical(pt, wl, ctx, workload, schedule):
    """b200pt_launch_PathTraceSamples: a frame of samples_per_frame = 5 passes as fused wavefronts of 1, 2, 4 or 5 passes in flight
    (CameraRays fused in, 1 or 2 path lanes) against the reference's loop of two launches per pass: accumulation buffer, denoiser
    AOVs and the G-buffer the last pass leaves are the same bits; two frames, so that batches start at a non-zero sample number."""
    W, H = (320, 200) if workload == "cornell" else (256, 160)
    cfg = getattr(wl, workload.replace("_ris", ""))(W, H)
    if workload.endswith("_ris"):
        cfg["options"] = dict(lss="ris", ess="none", ggx=0, later=3)       # RIS (the reference's default strategy) batches as well
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule=schedule)
    names = ("pixels", "denoiser_albedo", "denoiser_normals", "first_hit_prim_index", "shading_normals", "geometric_normals", "view_directions", "first_hits",
             "camera_ray_hit", "materials", "ray_volume_states", "pixel_active")
    results = {}
    for in_flight, lanes, bin_bits in ((0, 1, 0), (1, 1, 0), (2, 2, 3), (4, 1, 0), (5, 2, 5), (0, 2, 4)):
        ctx.set_ray_bin_bits(bin_bits)         # ray binning before traversal only reorders the queues: same bits
        g.set_samples_in_flight(in_flight)
        g.set_path_lanes(lanes)
        g.reset()
        g.resize(W, H)
        wl.apply_settings(g.render_settings, cfg["settings"])
        g.render_settings.samples_per_frame = 5
        assert g.renders_sample_batches == (in_flight > 0)
        for _ in range(2):
            g.render_frame()
        g.synchronize()
        assert g.render_settings.sample_number == 10
        results[(in_flight, lanes, bin_bits)] = {n: g.download(n).copy() for n in names}
    ctx.set_ray_bin_bits(-1)
    base = results[(0, 1, 0)]
    assert base["camera_ray_hit"].any() and np.isfinite(base["pixels"]).all() and base["pixels"].sum() > 0
    hit = base["camera_ray_hit"].astype(bool)
    for key, r in results.items():
        for n in names:
            a, b = base[n], r[n]
            if n in ("shading_normals", "geometric_normals", "first_hits", "materials", "ray_volume_states"):
                a, b = a[hit], b[hit]          # CameraRays leaves these untouched where the camera ray missed
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), f"{n} differs with {key[0]} passes in flight, {key[1]} lanes, ray bin bits {key[2]}"
    g.close()


def test_alpha_testing_statistical_parity(pt, wl, ctx, ref):
    """Row F, the any-hit filter with do_alpha_testing: a candidate hit survives iff rng() < alpha_opacity * texture alpha.
    The number drawn per candidate depends on the traversal order of the acceleration structure, so the streams of the two
    sides differ by construction; the images must agree statistically (block means), and must differ from the opaque render."""
    W, H, spp = 96, 64, 96
    cfg = wl.golden_scene("cornell_pbr", W, H, alpha_opacity=0.35)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    r.render(spp)
    for _ in range(spp):
        g.render_frame()
    g.synchronize()
    a = r.pixels().reshape(H, W, 3) / spp
    b = g.download("pixels").reshape(H, W, 3) / spp
    opaque_cfg = wl.golden_scene("cornell_pbr", W, H)
    o = wl.setup_gpu(pt, ctx, opaque_cfg, seed_schedule="cpu")
    for _ in range(spp):
        o.render_frame()
    o.synchronize()
    c = o.download("pixels").reshape(H, W, 3) / spp

    def blocks(img):
        return np.clip(img, 0.0, 10.0).reshape(H // 16, 16, W // 16, 16, 3).mean(axis=(1, 3, 4))
    ba, bb, bc = blocks(a), blocks(b), blocks(c)
    err = np.abs(ba - bb).mean() / ba.mean()             # same estimator, different random streams
    effect = np.abs(bc - bb).mean() / ba.mean()          # what the filter changes in this scene (sky-lit box: a few percent)
    assert err < 0.02, err
    assert effect > 0.03 and effect > 3.0 * err, (effect, err)
    assert abs(ba.mean() - bb.mean()) / ba.mean() < 0.02
    g.close()
    o.close()
