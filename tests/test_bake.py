"""Directional-albedo LUT bakers (include/b200pt_bake.h, csrc/b200pt_bake.cu) against the reference's own bake kernels
(src/Device/kernels/Baking/*.h compiled for the CPU: oracle/ref_bake.cpp; golden vectors tests/golden/bake_tables.npz made by
tools/make_bake_golden.py) and against the tables the reference ships (data/brdf_luts.npz, packed from its data/BRDFsData/*.hdr)."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
GOLDEN = os.path.join(ROOT, "tests", "golden", "bake_tables.npz")
NAN_SHARE_THIN_GLASS = 0.05
NAN_MASK_MISMATCH_THIN_GLASS = 0.02      # measured on B200: 0.13 % (small table), 0.16 % (full table: 2613 NaN texels here, 2615 in the reference)
KINDS = {"conductor": 0, "fresnel": 1, "glossy": 2, "glass": 3, "glass_inverse": 4, "thin_glass": 5}


@pytest.fixture(scope="module")
def bk():
    return __import__("importlib").import_module("hiprt-path-tracer_b200.baker")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


# ---- host logic + oracle (no GPU) ---------------------------------------------------------------------------------------------------------
def test_bake_plan_restates_the_reference_launch_arithmetic(bk, golden):
    """b200pt_bake_plan_for against GPUBakerKernel.cpp:82-83 as the oracle ran it (the plan stored with every golden table), the shipped sizes,
    and the divisor each kernel recomputes -- the Fresnel kernel's leaves the ior axis out (GGXFresnelDirectionalAlbedo.h:63-65)."""
    for key in golden.files:
        if key.endswith("_plan"):
            sx, sy, sz, count, ipk, launches = (int(v) for v in golden[key])
            kind = KINDS[key[len("small/"):-len("_plan")].rsplit("_", 1)[0]]
            p = bk.bake_plan(kind, sx, sy, sz, count)
            assert (p.iterations_per_kernel, p.nb_kernel_launch) == (ipk, launches), key
            assert p.texels == sx * sy * sz and 1 <= p.chunks <= 16
    # shipped resolutions, 65536 samples (GPUBakerConstants.h:14-31)
    p = bk.bake_plan(0, 128, 128, 1, 65536)
    assert (p.iterations_per_kernel, p.nb_kernel_launch, p.nb_samples) == (6103, 11, 67133)
    p = bk.bake_plan(1, 256, 256, 256, 65536)
    assert (p.iterations_per_kernel, p.nb_kernel_launch) == (5, 13108)
    assert p.nb_samples == 43 * 1525                      # ceil(65536 / floor(1e8 / 65536)) * 1525: not the 65540 samples a texel integrates
    p = bk.bake_plan(2, 128, 64, 128, 65536)
    assert (p.iterations_per_kernel, p.nb_kernel_launch, p.nb_samples) == (95, 690, 65550)
    p = bk.bake_plan(5, 32, 32, 96, 65536)
    assert (p.iterations_per_kernel, p.nb_kernel_launch, p.nb_samples) == (1017, 65, 66105)
    # chunks: explicit, automatic (small tables are split, large ones are not), clamped to the number of launches
    assert bk.bake_plan(0, 128, 128, 1, 65536, bk.BakeOptions(0, 0, 1)).chunks == 1
    assert bk.bake_plan(0, 128, 128, 1, 65536).chunks == 11                  # 11 launches: one per thread (8 threads would leave two idle)
    assert bk.bake_plan(5, 32, 32, 96, 65536).chunks == 3                   # 65 launches: 22 + 22 + 21
    assert bk.bake_plan(2, 128, 64, 128, 65536).chunks == 1
    assert bk.bake_plan(0, 16, 16, 1, 4096, bk.BakeOptions(0, 0, 8)).chunks == 1
    lib = bk._capi.lib()
    plan = bk.BakePlan()
    for bad in ((6, 16, 16, 1, 64), (0, 1, 16, 1, 64), (2, 16, 16, 1, 64), (0, 16, 16, 1, 0), (2, 2048, 2048, 1024, 64)):
        assert lib.b200pt_bake_plan_for(*[C.c_int(v) for v in bad], None, C.byref(plan)) == -1, bad
    assert lib.b200pt_bake_plan_for(C.c_int(0), C.c_int(16), C.c_int(16), C.c_int(1), C.c_int(64), C.byref(bk.BakeOptions(0, 0, 17)), C.byref(plan)) == -1


def test_reference_file_names(bk):
    assert bk.get_GGX_conductor_Ess_filename() == "GGX_Conductor_128x128.hdr"
    assert bk.get_glossy_dielectric_Ess_filename() == "Glossy_Ess_128x64x128.hdr"
    assert bk.get_GGX_glass_Ess_filename() == "GGX_Glass_Ess_256x16x128.hdr" and bk.get_GGX_glass_inv_Ess_filename() == "inv_GGX_Glass_Ess_256x16x128.hdr"
    assert bk.get_GGX_thin_glass_Ess_filename() == "GGX_Thin_Glass_Ess_32x32x96.hdr" and bk.get_GGX_fresnel_Ess_filename() == "GGX_Fresnel_256x256x256.hdr"


def test_oracle_bake_reproduces_golden_and_the_shipped_conductor_table(golden, pt):
    """The oracle pin: the reference's bake kernel run here gives the committed golden table again, and the golden full-size conductor table
    IS the table the reference ships (GGX_Conductor_128x128.hdr through its own loader, y flipped on load) to the 8-bit RGBE mantissa."""
    from oracle import ref
    if not os.path.exists(ref.bake_lib_path("vndf")):
        pytest.skip("oracle/_ref bake library is not built")
    table, ipk, launches = ref.bake(5, 12, 8, 8, 4096)
    assert np.array_equal(table, golden["small/thin_glass_vndf"], equal_nan=True)
    shipped = pt.load_brdf_luts()["conductor"][::-1]
    d = np.abs(golden["full/conductor"] - shipped)
    assert d.max() < 5e-3 and d.mean() < 2.5e-3, (d.max(), d.mean())


def test_host_hdr_writer_is_read_back_by_the_reference_loader(pt, tmp_path):
    """The C++ host layer's slice writer (GPUBaker.cpp: write_image_hdr_1channel) on a table-like array with the awkward values in it: what it
    writes decodes -- with the reference's own loader (Image32Bit::read_image_hdr over stb_image) where the oracle is built, and with a plain
    RGBE decoder -- to the input within the 8-bit mantissa (1 / 256 relative, rounded down); zeros, denormal-small and non-finite texels give 0."""
    host = pt.host()
    rng = np.random.default_rng(3)
    table = (rng.random((24, 40)) ** 3).astype(np.float32)
    table[0, :6] = [0.0, 1.0, 1.0e-35, np.nan, np.inf, 0.999999]
    table[1, 0] = 7.5
    path = str(tmp_path / "slice.hdr")
    pt._capi.check_host(host.b200pt_host_write_hdr_1channel(path.encode(), table.ctypes.data_as(C.c_void_p), 40, 24), "b200pt_host_write_hdr_1channel")
    assert host.b200pt_host_write_hdr_1channel(str(tmp_path / "missing" / "x.hdr").encode(), table.ctypes.data_as(C.c_void_p), 40, 24) == -1
    want = np.where(np.isfinite(table) & (table >= 1.0e-32), table, 0.0).astype(np.float32)
    data = open(path, "rb").read()
    head, _, rest = data.partition(b"\n\n")
    assert head.startswith(b"#?RADIANCE") and b"FORMAT=32-bit_rle_rgbe" in head and rest.startswith(b"-Y 24 +X 40\n")
    decoded = [_read_flat_hdr_file(path)]
    from oracle import ref
    if ref.available("mis"):
        lib = ref.load("mis")
        got = np.empty((24, 40), np.float32)
        w, h = C.c_int(), C.c_int()
        assert lib.ref_load_hdr(path.encode(), 1, 0, got.ctypes.data_as(C.c_void_p), got.size, C.byref(w), C.byref(h)) == 0 and (w.value, h.value) == (40, 24)
        decoded.append(got)
    for got in decoded:
        assert got.shape == want.shape
        assert np.all(got <= want * (1.0 + 1e-6)) and np.all(want - got <= want / 128.0), float(np.abs(got - want).max())
        assert got[0, 0] == 0.0 and got[0, 2] == 0.0 and got[0, 3] == 0.0 and got[0, 4] == 0.0 and got[0, 1] == 1.0


def _read_flat_hdr_file(path):
    io = __import__("importlib").import_module("hiprt-path-tracer_b200.image_io")
    data = open(path, "rb").read()
    _, _, rest = data.partition(b"\n\n")
    dims, _, body = rest.partition(b"\n")
    _, h, _, w = dims.split()
    return io.rgbe_to_float(np.frombuffer(body, np.uint8).reshape(int(h), int(w), 4))[..., 0]


# ---- GPU ------------------------------------------------------------------------------------------------------------------------------------
def _compare(got, want, what, max_abs=1e-3, mean_abs=1e-5, nan_mismatch=0.0):
    """Texels finite on both sides to the tolerance; the thin-glass kernel of the reference ITSELF yields NaN texels (ior -> 1: the generalised
    half-vector term HoL + HoV / eta of GGXThinGlassDirectionalAlbedo.h:177-179 rounds to exactly 0 for one sample in ~1e5 and inf / inf poisons the
    texel's sum: 2.7 % of the full-size table, most of its ior slices 0-2), so NaN is compared as a mask. Whether one of a texel's 1e5 samples
    hits exactly 0 depends on the last ulp of sinf / cosf, where CUDA and glibc differ for about one sample in five: the masks agree on most
    texels, not on all (NAN_MASK_MISMATCH_THIN_GLASS of the table at most)."""
    assert got.shape == want.shape, what
    fg, fw = np.isfinite(got), np.isfinite(want)
    mismatch = float((fg != fw).mean())
    both = fg & fw
    d = np.abs(got[both] - want[both])
    stats = (float(d.max()), float(d.mean()), int((~fg).sum()), int((~fw).sum()), mismatch)
    assert d.max() < max_abs and d.mean() < mean_abs and mismatch <= nan_mismatch, (what, stats)
    return stats


def _bake_small(bk, baker, name, sx, sy, sz, count):
    if name == "conductor":
        return baker.bake_ggx_conductor_directional_albedo(bk.GGXConductorDirectionalAlbedoSettings(sx, sy, count))[None]
    if name == "fresnel":
        return baker.bake_ggx_fresnel_directional_albedo(bk.GGXFresnelDirectionalAlbedoSettings(sx, sy, sz, count))
    if name == "glossy":
        return baker.bake_glossy_dielectric_directional_albedo(bk.GlossyDielectricDirectionalAlbedoSettings(sx, sy, sz, count))
    if name in ("glass", "glass_inverse"):
        return baker.bake_ggx_glass_directional_albedo(bk.GGXGlassDirectionalAlbedoSettings(sx, sy, sz, count))[0 if name == "glass" else 1]
    return baker.bake_ggx_thin_glass_directional_albedo(bk.GGXThinGlassDirectionalAlbedoSettings(sx, sy, sz, count))


@pytest.mark.gpu
def test_bake_kernels_match_the_reference_kernels_on_small_tables(bk, ctx, golden):
    """Every kernel, both GGX sampling routines, same seeds and samples as the oracle run. With chunks = 1 the float sum is the reference's, so
    what is left is libm (sinf/cosf/powf/acosf ulps flipping a rare sample): 1e-3 of the table's range at worst, 1e-5 on average. Automatic
    chunking regroups the sum of the same samples."""
    worst = {}
    for key in golden.files:
        if not key.startswith("small/") or key.endswith("_plan"):
            continue
        name, sampler = key[len("small/"):].rsplit("_", 1)
        sx, sy, sz, count, _, _ = (int(v) for v in golden[key + "_plan"])
        want = golden[key]
        exact = bk.GPUBaker(ctx, ggx_sample_function=1 if sampler == "caps" else 0, chunks=1)
        got = _bake_small(bk, exact, name, sx, sy, sz, count)
        worst[key] = _compare(got, want, key, nan_mismatch=NAN_MASK_MISMATCH_THIN_GLASS if name == "thin_glass" else 0.0)
        auto = bk.GPUBaker(ctx, ggx_sample_function=1 if sampler == "caps" else 0)
        split = _bake_small(bk, auto, name, sx, sy, sz, count)
        assert np.array_equal(np.isfinite(split), np.isfinite(got)) and np.nanmax(np.abs(split - got)) < 2e-5, key
    print("bake parity (max abs, mean abs, NaN texels here, NaN texels in the reference, mask mismatch):", worst)


@pytest.mark.gpu
def test_full_size_tables_match_the_reference_slices_and_the_shipped_luts(bk, ctx, golden, pt, tmp_path):
    """The tables the reference ships, regenerated at their shipped resolution and sample count (GPUBakerConstants.h:14-31; 1.5e11 samples):
    (1) the ior slices the oracle integrated (tests/golden) to the libm tolerance above, (2) every texel against data/brdf_luts.npz, which went
    through an 8-bit-mantissa Radiance file (1/256 relative) -- a different run of the same estimator would also differ by its Monte-Carlo noise,
    this one does not (same seeds). Files are written with the reference's names."""
    luts = pt.load_brdf_luts()
    # chunks = 1: a texel's float sum in the reference's order (the 3D tables but the thin-glass one get that by default: they fill the GPU)
    baker = bk.GPUBaker(ctx, chunks=1, output_directory=str(tmp_path))
    tables = {"conductor": baker.bake_ggx_conductor_directional_albedo(output_filename=bk.get_GGX_conductor_Ess_filename())[None],
              "glossy": baker.bake_glossy_dielectric_directional_albedo(output_filename=bk.get_glossy_dielectric_Ess_filename()),
              "thin_glass": baker.bake_ggx_thin_glass_directional_albedo(output_filename=bk.get_GGX_thin_glass_Ess_filename())}
    tables["glass"], tables["glass_inverse"] = baker.bake_ggx_glass_directional_albedo(output_filename=bk.get_GGX_glass_Ess_filename())
    assert baker.is_ggx_glass_directional_albedo_bake_complete() and not baker.is_ggx_fresnel_directional_albedo_bake_complete()
    # the default plan splits the two small tables over 11 and 3 threads per texel: same samples, the sum regrouped. 67 k additions of ~1.5e-5 to a
    # running sum near 1 carry ~1e-4 of rounding in the reference's sequential order itself, so the regrouped sum differs from it by that much
    split = bk.GPUBaker(ctx)
    assert bk.bake_plan(0, 128, 128, 1, 65536).chunks == 11 and bk.bake_plan(5, 32, 32, 96, 65536).chunks == 3
    regrouped = {"conductor": split.bake_ggx_conductor_directional_albedo()[None], "thin_glass": split.bake_ggx_thin_glass_directional_albedo()}
    for name, table in regrouped.items():
        assert np.array_equal(np.isfinite(table), np.isfinite(tables[name])), name
        d = np.abs(table - tables[name])[np.isfinite(table)]
        assert d.max() < 1.5e-3 and d.mean() < 3e-4, (name, d.max(), d.mean())
    report = {}
    for key in golden.files:
        if key.startswith("slice/"):
            name, z = key[len("slice/"):].rsplit("_z", 1)
            report[key] = _compare(tables[name][int(z)], golden[key], key, nan_mismatch=NAN_MASK_MISMATCH_THIN_GLASS if name == "thin_glass" else 0.0)
    report["full/conductor"] = _compare(tables["conductor"][0], golden["full/conductor"], "conductor")
    report["full/thin_glass"] = _compare(tables["thin_glass"], golden["full/thin_glass"], "thin_glass", nan_mismatch=NAN_MASK_MISMATCH_THIN_GLASS)
    for name, table in tables.items():
        shipped = luts[name][None, ::-1] if name == "conductor" else luts[name][:, ::-1]          # the loader flips y (CPURenderer.cpp:93-132)
        assert shipped.shape == table.shape, name
        if name in ("glass", "glass_inverse"):
            # the shipped glass tables predate the current kernel's cos(theta_o)^2.5 axis (the oracle slices differ from them in the same way:
            # tools/make_bake_golden.py output vs data/brdf_luts.npz); only the grazing column x = 0 is the same point of the domain in both
            table, shipped = table[:, :, :1], shipped[:, :, :1]
        fin = np.isfinite(table)
        d = np.abs(table[fin] - shipped[fin])
        report["shipped/" + name] = (float(d.max()), float(d.mean()), float(np.quantile(d, 0.999)), int((~fin).sum()))
        # a shipped table is another run of the estimator (other seeds or an older build: Monte-Carlo noise) behind an 8-bit mantissa; the oracle
        # slices differ from it by up to 0.0027 mean / 0.015 at the 99.9th percentile. NaN texels: the thin-glass table only (see _compare)
        assert (~fin).mean() <= (NAN_SHARE_THIN_GLASS if name == "thin_glass" else 0.0), (name, report["shipped/" + name])
        # (the shipped glass tables also hold 0 in their roughness-0 row where the current kernel integrates 0.005-0.23: mean only)
        assert d.mean() < 4e-3 and (name.startswith("glass") or np.quantile(d, 0.999) < 2e-2), (name, report["shipped/" + name])
    ms = baker.last_kernel_ms
    samples = {"ggx_conductor": 128 * 128 * 67133, "glossy_dielectric": 128 * 64 * 128 * 65550, "ggx_glass": 2 * 256 * 16 * 128 * 65550, "ggx_thin_glass": 32 * 32 * 96 * 66105}
    print("full-size bakes, one thread per texel:", {k: f"{ms[k]:.1f} ms, {samples[k] / ms[k] / 1e6:.1f} Gsamples/s" for k in samples},
          "default plan:", {k: f"{split.last_kernel_ms[k]:.1f} ms, {samples[k] / split.last_kernel_ms[k] / 1e6:.1f} Gsamples/s" for k in split.last_kernel_ms}, report)
    # files: GPUBakerKernel.cpp:107-126 writes '<slice><name>' per ior slice
    for name in ("GGX_Conductor_128x128.hdr", "0Glossy_Ess_128x64x128.hdr", "127Glossy_Ess_128x64x128.hdr", "95GGX_Thin_Glass_Ess_32x32x96.hdr",
                 "0GGX_Glass_Ess_256x16x128.hdr", "127inv_GGX_Glass_Ess_256x16x128.hdr"):
        assert os.path.getsize(os.path.join(str(tmp_path), name)) > 1000, name


@pytest.mark.gpu
def test_bake_argument_errors(bk, ctx):
    lib = bk._capi.lib()
    s = bk.GGXConductorDirectionalAlbedoSettings(16, 16, 64)
    assert lib.b200pt_bake_ggx_conductor_directional_albedo(ctx.handle, C.byref(s), None, None, None) == -1                  # no output
    buf = ctx.malloc(4 * 256)
    try:
        assert lib.b200pt_bake_ggx_conductor_directional_albedo(ctx.handle, C.byref(s), C.byref(bk.BakeOptions(2, 0, 0)), buf, None) == -1   # unknown GGX sampler
        assert lib.b200pt_bake_ggx_conductor_directional_albedo(None, C.byref(s), None, buf, None) == -1
        g = bk.GGXGlassDirectionalAlbedoSettings(8, 8, 8, 64)
        assert lib.b200pt_bake_ggx_glass_directional_albedo(ctx.handle, C.byref(g), None, None, None, None) == -1           # neither table wanted
    finally:
        ctx.free(buf)


def _read_flat_hdr(path):
    """Reader for the flat (not run-length encoded) Radiance files the bakers write."""
    io = __import__("importlib").import_module("hiprt-path-tracer_b200.image_io")
    data = open(path, "rb").read()
    head, _, rest = data.partition(b"\n\n")
    assert head.startswith(b"#?RADIANCE") and b"32-bit_rle_rgbe" in head
    dims, _, body = rest.partition(b"\n")
    _, h, _, w = dims.split()
    rgbe = np.frombuffer(body, np.uint8).reshape(int(h), int(w), 4)
    return io.rgbe_to_float(rgbe)[..., 0]


@pytest.mark.gpu
def test_host_layer_gpubaker_is_asynchronous_and_writes_the_reference_files(bk, pt, ctx, golden, tmp_path):
    """b200pt_host::GPUBaker (C++, hiprt-path-tracer_b200/host/GPUBaker.cpp) through its C handles: bake_*() returns while the table is being
    integrated, is_*_complete() goes from false to true (GPUBakerKernel's flag, src/Renderer/Baker/GPUBakerKernel.cpp:24,130), the tables equal the
    reference kernels' (golden), and the files carry the reference's names and decode -- with the reference's own loader where the oracle is
    built -- to the table within the 8-bit RGBE mantissa."""
    host = pt.host()
    baker = C.c_void_p()
    pt._capi.check_host(host.b200pt_baker_create(ctx.handle, C.byref(baker)), "b200pt_baker_create")
    try:
        assert host.b200pt_baker_is_complete(baker, 0) == 1                                   # never started: complete, as in the reference
        pt._capi.check_host(host.b200pt_baker_set_options(baker, 0, 0, 1, str(tmp_path).encode()), "b200pt_baker_set_options")
        assert host.b200pt_baker_bake(baker, 4, 12, 8, 8, 4096, None) == -1                   # the exiting glass table is not baked on its own
        assert host.b200pt_baker_bake(baker, 0, 1, 16, 1, 4096, None) == -1                   # sizes are validated before a worker starts
        # the big one first: 2e9 samples keep the GPU busy for tens of milliseconds, long enough to see the flag down
        pt._capi.check_host(host.b200pt_baker_bake(baker, 2, 64, 32, 16, 65536, b"Glossy_Ess_64x32x16.hdr"), "b200pt_baker_bake")
        in_flight = host.b200pt_baker_is_complete(baker, 2) == 0
        pt._capi.check_host(host.b200pt_baker_bake(baker, 0, 16, 16, 1, 4096, bk.get_GGX_conductor_Ess_filename(16, 16).encode()), "b200pt_baker_bake")
        pt._capi.check_host(host.b200pt_baker_bake(baker, 3, 12, 8, 8, 4096, bk.get_GGX_glass_Ess_filename(12, 8, 8).encode()), "b200pt_baker_bake")
        pt._capi.check_host(host.b200pt_baker_bake(baker, 5, 12, 8, 8, 4096, bk.get_GGX_thin_glass_Ess_filename(12, 8, 8).encode()), "b200pt_baker_bake")
        pt._capi.check_host(host.b200pt_baker_wait(baker), "b200pt_baker_wait")
        assert in_flight and all(host.b200pt_baker_is_complete(baker, k) == 1 for k in (0, 2, 3, 4, 5))
        tables = {}
        for kind, shape, key in ((0, (1, 16, 16), "small/conductor_vndf"), (3, (8, 8, 12), "small/glass_vndf"), (4, (8, 8, 12), "small/glass_inverse_vndf"),
                                 (5, (8, 8, 12), "small/thin_glass_vndf")):
            out = np.empty(shape, np.float32)
            ms = C.c_float()
            n = host.b200pt_baker_last_table(baker, kind, out.ctypes.data_as(C.c_void_p), C.c_longlong(out.size), C.byref(ms))
            assert n == out.size and ms.value > 0.0, (kind, n)
            _compare(out, golden[key], key, nan_mismatch=NAN_MASK_MISMATCH_THIN_GLASS if kind == 5 else 0.0)
            tables[kind] = out
        assert host.b200pt_baker_last_table(baker, 0, None, C.c_longlong(0), None) == -256    # capacity too small: the size, negated
        names = ["GGX_Conductor_16x16.hdr"] + [f"{z}{n}" for z in (0, 7) for n in ("GGX_Glass_Ess_12x8x8.hdr", "inv_GGX_Glass_Ess_12x8x8.hdr", "GGX_Thin_Glass_Ess_12x8x8.hdr")]
        names += ["0Glossy_Ess_64x32x16.hdr", "15Glossy_Ess_64x32x16.hdr"]
        for name in names:
            assert os.path.exists(os.path.join(str(tmp_path), name)), name
        for name, table in (("GGX_Conductor_16x16.hdr", tables[0][0]), ("7GGX_Glass_Ess_12x8x8.hdr", tables[3][7]), ("3inv_GGX_Glass_Ess_12x8x8.hdr", tables[4][3])):
            path = os.path.join(str(tmp_path), name)
            decoded = _read_flat_hdr(path)
            assert decoded.shape == table.shape and np.all(np.abs(decoded - table) <= table / 128.0 + 1e-30), name
            from oracle import ref
            if ref.available("mis"):
                # Image32Bit::read_image_hdr of the reference (stb_image), unflipped, first channel
                lib = ref.load("mis")
                got = np.empty(table.shape, np.float32)
                w, h = C.c_int(), C.c_int()
                assert lib.ref_load_hdr(path.encode(), 1, 0, got.ctypes.data_as(C.c_void_p), got.size, C.byref(w), C.byref(h)) == 0 and (w.value, h.value) == table.shape[::-1]
                assert np.all(np.abs(got - table) <= table / 128.0 + 1e-30), name
    finally:
        host.b200pt_baker_destroy(baker)
