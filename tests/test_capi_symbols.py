"""The C-ABI libraries load and export every symbol include/*.h declares; without a GPU they refuse to compute."""
import ctypes
import os
import re

import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    return sorted(set(re.findall(r"B200PT_API\s+[\w\s\*]+?\b(b200pt_\w+)\s*\(", text)))


def test_every_declared_symbol_is_exported(pt):
    lib, host = pt.lib(), pt.host()
    core, hostsyms = _declared("b200pt.h"), _declared("b200pt_host.h")
    assert len(core) >= 40 and len(hostsyms) >= 20
    for s in core:
        assert hasattr(lib, s), s
    for s in hostsyms:
        assert hasattr(host, s), s
    capi = __import__("importlib").import_module("hiprt-path-tracer_b200._capi")
    assert sorted(capi.API_SYMBOLS) == core
    assert sorted(capi.HOST_API_SYMBOLS) == hostsyms
    commsyms = _declared("b200pt_comm.h")            # the multi-GPU plumbing is part of libb200pt.so (NCCL itself is bound at run time)
    assert len(commsyms) == 14 and sorted(capi.COMM_API_SYMBOLS) == commsyms
    for s in commsyms:
        assert hasattr(lib, s), s
    bakesyms = _declared("b200pt_bake.h")            # the LUT bakers
    assert len(bakesyms) == 6 and sorted(capi.BAKE_API_SYMBOLS) == bakesyms
    for s in bakesyms:
        assert hasattr(lib, s), s


def test_no_cpu_fallback(pt):
    lib = pt.lib()
    if lib.b200pt_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pt.B200ptError, match="no CPU fallback"):
        pt.Context(0)
    # argument validation does not need a device
    assert lib.b200pt_ctx_create(0, None) == -1


def test_variant_list_matches_header(pt):
    text = open(os.path.join(ROOT, "hiprt-path-tracer_b200", "csrc", "variants.h")).read()
    variants = re.findall(r"M\((\d+),\s*(\d+),\s*(\d+),\s*(\d+)\)", text.split("PT_FOR_EACH_VARIANT(M)")[1])
    assert ("3", "0", "3", "0") in variants and ("3", "2", "3", "0") in variants


def test_comm_tile_plan_matches_the_python_plan(pt):
    """b200pt_comm_tile (C ABI, needs neither GPU nor NCCL) against multigpu.row_tiles / column_tiles, which the gloo tests cover."""
    mg = __import__("importlib").import_module("hiprt-path-tracer_b200.multigpu")
    lib = pt.lib()
    for width, height in ((1920, 1080), (3840, 2160), (96, 64), (17, 9), (8, 8)):
        for world in (1, 2, 3, 4, 8):
            rows, cols = mg.row_tiles(height, world), mg.column_tiles(width, world)
            for rank in range(world):
                assert mg.Comm.tile(lib, mg.Comm.ROWS, width, height, world, rank) == (0, rows[rank][0], width, rows[rank][1])
                assert mg.Comm.tile(lib, mg.Comm.COLUMNS, width, height, world, rank) == (cols[rank][0], 0, cols[rank][1], height)
    t = pt.structs.Tile()
    assert lib.b200pt_comm_tile(0, 64, 64, 2, 2, ctypes.byref(t)) == -1          # rank out of range
    assert lib.b200pt_comm_restir_halo_rows(16, 4, 0) == 20


def test_comm_balanced_bounds_equalise_a_known_cost_profile(pt):
    """b200pt_comm_balanced_bounds (C ABI, no GPU): against a per-column cost profile (a cheap 'sky' third, an expensive middle, a ramp) the
    plan converges in a few iterations to bands whose costs differ by no more than an alignment block; boundaries stay aligned, monotonic, and cover the image.
    Degenerate inputs: zero costs keep the plan, a world larger than the number of blocks leaves empty bands at the end."""
    import numpy as np
    mg = __import__("importlib").import_module("hiprt-path-tracer_b200.multigpu")
    lib = pt.lib()
    W = 1920
    x = np.arange(W)
    column_cost = np.where(x < 600, 1.0, np.where(x < 1300, 4.0, 1.0 + 3.0 * (x - 1300) / 620.0))
    for world in (2, 3, 8):
        bounds = [t[0] for t in mg.column_tiles(W, world)] + [W]
        first = None
        for _ in range(4):
            costs = [float(column_cost[bounds[r]:bounds[r + 1]].sum()) for r in range(world)]
            first = first or max(costs) / (sum(costs) / world)
            bounds = mg.Comm.balanced_bounds(lib, bounds, costs, W)
            assert bounds[0] == 0 and bounds[-1] == W and all(b % 8 == 0 for b in bounds[:-1])
            assert all(bounds[r + 1] - bounds[r] >= 8 for r in range(world))
        costs = [float(column_cost[bounds[r]:bounds[r + 1]].sum()) for r in range(world)]
        assert max(costs) / (sum(costs) / world) < 1.08 <= first, (world, first, costs)      # one 8-column block of the expensive part is 5 % of a band at 8 ranks
    assert mg.Comm.balanced_bounds(lib, [0, 8, 16], [0.0, 0.0], 16) == [0, 8, 16]
    assert mg.Comm.balanced_bounds(lib, [0, 8, 16, 16], [1.0, 3.0, 0.0], 16) == [0, 8, 16, 16]        # two blocks, three ranks: the last band stays empty
    out = (ctypes.c_int * 3)()
    assert lib.b200pt_comm_balanced_bounds((ctypes.c_int * 3)(0, 8, 15), (ctypes.c_double * 2)(1.0, 1.0), 2, 16, 8, out) == -1      # bounds must end at the extent
