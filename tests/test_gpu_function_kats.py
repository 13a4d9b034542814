"""Per-function known-answer pins: the CUDA device functions of the shading path, called through the C ABI
(b200pt_debug_*), against tests/golden/function_kats.npz, the reference's own CPU kernels' answers on the G-buffer inputs of its own render
(tools/make_function_kats.py; the non-GPU test test_oracle_golden.py::test_function_kats_reproduce keeps the file honest).
Arithmetic is the reference's, in its order, with -fmad=false: values agree to a few ulp except where a glibc / CUDA sinf-cosf-powf ulp
flips a branch, so the bar is: relative 2e-4 (absolute 1e-6) on >= 99.5 % of the items, generator states equal on >= 99.5 %."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu


def _close_fraction(a, b, rtol=2e-4, atol=1e-6):
    ok = np.isclose(a, b, rtol=rtol, atol=atol) | (np.isnan(a) & np.isnan(b))
    return float(ok.all(axis=1).mean())


@pytest.fixture(scope="module")
def kat_gpu(pt, wl, ctx):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_function_kats as mk
    cfg = mk.kat_config()
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    g.render_frame()                      # fills the render data's device pointers (LUTs, envmap, alias table)
    g.synchronize()
    gold = np.load(os.path.join(GOLDEN, "function_kats.npz"))
    yield g, gold
    g.close()


def _run(ctx, fn, n, out_cols, *device_args):
    d_out = ctx.malloc(4 * n * out_cols)
    fn(d_out)
    ctx.sync()
    out = ctx.download(d_out, (n, out_cols), np.float32)
    ctx.free(d_out)
    return out


def test_bsdf_eval_and_sample_kats(pt, ctx, kat_gpu):
    g, gold = kat_gpu
    lib, rd = ctx.lib, C.byref(g.render_data)
    n = len(gold["seeds"])
    d_mats = ctx.upload(gold["materials"])
    d_eval_vecs, d_sample_vecs = ctx.upload(gold["eval_vecs"]), ctx.upload(gold["sample_vecs"])
    try:
        d_vs = ctx.upload(gold["volume_states"])
        ev = _run(ctx, lambda d_out: pt._capi.check(lib.b200pt_debug_bsdf_eval(ctx.handle, rd, d_mats, d_vs, d_eval_vecs, C.c_longlong(n), d_out), "b200pt_debug_bsdf_eval"), n, 4)
        ctx.free(d_vs)
        frac = _close_fraction(ev, gold["eval"])
        assert frac >= 0.995, f"bsdf_eval: {frac:.4f} of {n} items within tolerance"
        for sampler, key in ((0, "vndf"), (1, "caps")):
            d_vs, d_seeds = ctx.upload(gold["volume_states"]), ctx.upload(gold["seeds"])
            sa = _run(ctx, lambda d_out: pt._capi.check(lib.b200pt_debug_bsdf_sample(ctx.handle, rd, d_mats, d_vs, d_sample_vecs, d_seeds, C.c_longlong(n), C.c_int(sampler), d_out),
                                                        "b200pt_debug_bsdf_sample"), n, 7)
            seeds = ctx.download(d_seeds, (n,), np.uint32)
            ctx.free(d_vs)
            ctx.free(d_seeds)
            want = gold[f"sample_{key}"]
            live = want[:, 3] > 0                      # a sample with pdf 0 carries no direction (the reference leaves it unset)
            frac = _close_fraction(sa[live], want[live])
            assert frac >= 0.995, f"bsdf_sample[{key}]: {frac:.4f} of {int(live.sum())} items within tolerance"
            assert ((sa[:, 3] > 0) == live).mean() >= 0.995
            assert (seeds == gold[f"sample_{key}_seeds"]).mean() >= 0.995, f"bsdf_sample[{key}]: generator states after sampling differ"
        assert (gold["sample_vndf"] != gold["sample_caps"]).any(axis=1).sum() > 1000      # the two samplers are told apart by the vectors
    finally:
        for p in (d_mats, d_eval_vecs, d_sample_vecs):
            ctx.free(p)


def test_envmap_sample_and_thin_film_kats(pt, ctx, kat_gpu):
    g, gold = kat_gpu
    lib, rd = ctx.lib, C.byref(g.render_data)
    n = len(gold["seeds"])
    d_seeds = ctx.upload(gold["seeds"])
    env = _run(ctx, lambda d_out: pt._capi.check(lib.b200pt_debug_envmap_sample(ctx.handle, rd, d_seeds, C.c_longlong(n), d_out), "b200pt_debug_envmap_sample"), n, 8)
    seeds = ctx.download(d_seeds, (n,), np.uint32)
    ctx.free(d_seeds)
    assert np.array_equal(seeds, gold["envmap_seeds"])                # integer work: the alias-table draw is exact
    frac = _close_fraction(env, gold["envmap"])
    assert frac >= 0.995, f"envmap_sample: {frac:.4f}"
    d_mats, d_args = ctx.upload(gold["materials"]), ctx.upload(gold["film_args"])
    film = _run(ctx, lambda d_out: pt._capi.check(lib.b200pt_debug_thin_film_fresnel(ctx.handle, d_mats, d_args, C.c_longlong(n), d_out), "b200pt_debug_thin_film_fresnel"), n, 3)
    ctx.free(d_mats)
    ctx.free(d_args)
    frac = _close_fraction(film, gold["film"], rtol=5e-4, atol=1e-5)
    assert frac >= 0.995, f"thin_film_fresnel: {frac:.4f}"
