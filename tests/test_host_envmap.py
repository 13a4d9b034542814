"""Host logic around the environment map (no GPU): the alias table the host layer builds must equal, entry for entry, the
table the reference's Image32Bit::compute_alias_table builds (src/Image/Image.cpp:579-659), and sampling it must
reproduce the luminance distribution."""
import ctypes as C
import importlib

import numpy as np


def _host_table(pt, texels):
    host = pt.host()
    h, w, ch = texels.shape
    probas, alias, total = np.empty(w * h, np.float32), np.empty(w * h, np.int32), C.c_float()
    rc = host.b200pt_host_compute_alias_table(texels.ctypes.data_as(C.c_void_p), C.c_int(w), C.c_int(h), C.c_int(ch),
                                              probas.ctypes.data_as(C.c_void_p), alias.ctypes.data_as(C.c_void_p), C.byref(total))
    assert rc == 0
    return probas, alias, total.value


def test_alias_table_equals_the_reference_builder(pt, ref):
    procedural = importlib.import_module("hiprt-path-tracer_b200.procedural")
    for (w, h, ch, seed) in ((256, 128, 4, 1234), (64, 32, 3, 7), (33, 17, 1, 9)):
        env = np.ascontiguousarray(procedural.procedural_envmap(w, h, seed)[..., :ch])
        r = ref.RefRenderer("mis")
        ref_probas, ref_alias, ref_total = r.set_envmap(env)
        probas, alias, total = _host_table(pt, env)
        assert total == ref_total
        assert np.array_equal(probas, ref_probas)
        # entries with probability 1 never read their alias
        used = probas < 1.0
        assert np.array_equal(alias[used], ref_alias[used])
        r.close()


def test_alias_table_reproduces_the_luminance_distribution(pt):
    rng = np.random.default_rng(0)
    env = rng.uniform(0.0, 1.0, (16, 32, 3)).astype(np.float32) ** 4
    env[3, 5] = 200.0    # a "sun"
    probas, alias, total = _host_table(pt, env)
    lum = (env.reshape(-1, 3).astype(np.float64) * np.array([0.3086, 0.6094, 0.0820], np.float32).astype(np.float64)).sum(axis=1)
    n = lum.size
    assert abs(total - lum.sum()) / lum.sum() < 1e-5
    # exact distribution implied by the table: P(i) = (probas[i] + sum over j aliasing to i of (1 - probas[j])) / n
    p = probas.astype(np.float64).copy()
    np.add.at(p, alias[probas < 1.0], 1.0 - probas[probas < 1.0].astype(np.float64))
    assert np.allclose(p / n, lum / lum.sum(), atol=1e-6)
    assert (probas >= 0).all() and (probas <= 1).all() and (alias >= 0).all() and (alias < n).all()


def test_single_texel_and_uniform_maps(pt):
    probas, alias, total = _host_table(pt, np.full((1, 1, 3), 2.0, np.float32))
    assert probas.tolist() == [1.0] and abs(total - 2.0) < 1e-6
    probas, alias, total = _host_table(pt, np.ones((4, 8, 4), np.float32))
    assert (probas == 1.0).all()
