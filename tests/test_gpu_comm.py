"""include/b200pt_comm.h on one GPU: a single-rank NCCL communicator runs every collective of the multi-GPU path through the C ABI
(the N > 1 behaviour of the same plan is covered on CPU by tests/test_multigpu_gloo.py and measured by bench.py --gpus N)."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_single_rank_communicator_runs_every_collective(pt, ctx):
    mg = importlib.import_module("hiprt-path-tracer_b200.multigpu")
    ident = mg.Comm.unique_id(ctx.lib)
    assert len(ident) == 128
    comm = mg.Comm(ctx, 0, 1, ident)
    try:
        W, H = 64, 40
        rng = np.random.default_rng(1)
        image = rng.random((H, W, 3), dtype=np.float32)
        d_image = ctx.upload(image)
        for sharding in (mg.Comm.ROWS, mg.Comm.COLUMNS):
            comm.gather_tiles(d_image.value, W, H, 12, sharding)
            ctx.sync()
            assert np.array_equal(ctx.download(d_image, (H, W, 3), np.float32), image)      # one rank owns the whole frame
        new_bounds, costs = comm.rebalance_columns([0, W], 3.5, W)         # cost-balanced plan: with one rank it is the whole frame, whatever the cost
        assert new_bounds == [0, W] and costs == [3.5]
        comm.gather_tiles(d_image.value, W, H, 12, mg.Comm.COLUMNS)
        ctx.sync()
        assert np.array_equal(ctx.download(d_image, (H, W, 3), np.float32), image)
        comm.set_column_bounds(None)
        d_sum = ctx.malloc(image.nbytes)
        comm.allreduce_sum(d_image.value, d_sum.value, image.size, "f32")
        counts = rng.integers(0, 100, (H, W)).astype(np.int32)
        d_counts = ctx.upload(counts)
        comm.allreduce_sum(d_counts.value, d_counts.value, counts.size, "i32")                  # in place is allowed
        reservoirs = rng.integers(0, 256, (H, W * 36), dtype=np.uint8)
        d_res = ctx.upload(reservoirs)
        comm.exchange_halo(d_res.value, H, W * 36, 20)
        ctx.sync()
        assert np.array_equal(ctx.download(d_sum, (H, W, 3), np.float32), image)
        assert np.array_equal(ctx.download(d_counts, (H, W), np.int32), counts)
        assert np.array_equal(ctx.download(d_res, (H, W * 36), np.uint8), reservoirs)
        assert comm.allreduce_host([1.5, -2.0], "max") == [1.5, -2.0] and comm.allreduce_host([3.0], "sum") == [3.0]
        comm.barrier()
        for p in (d_image, d_sum, d_counts, d_res):
            ctx.free(p)
    finally:
        comm.close()
