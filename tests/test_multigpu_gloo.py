"""Host-side logic of the multi-GPU path on CPU: tile plan, ReSTIR halo plan, and the three collectives over gloo with
world_size 2 and 3 (the same functions run over NCCL on the GPUs, bench.py)."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

mg = importlib.import_module("hiprt-path-tracer_b200.multigpu")


def test_row_tiles_cover_the_image_in_8_row_blocks():
    for height in (0, 1, 7, 8, 96, 1080, 2160):
        for world in (1, 2, 3, 4, 8):
            tiles = mg.row_tiles(height, world)
            assert len(tiles) == world
            assert tiles[0][0] == 0 and tiles[-1][1] == height
            for (a0, a1), (b0, b1) in zip(tiles[:-1], tiles[1:]):
                assert a1 == b0 and a0 <= a1
                assert a1 % 8 == 0 or a1 == height
    assert mg.row_tiles(1080, 8) == [(0, 128), (128, 264), (264, 400), (400, 536), (536, 672), (672, 808), (808, 944), (944, 1080)]
    with pytest.raises(ValueError):
        mg.row_tiles(100, 0)


def test_halo_plan_is_symmetric_and_reaches_far_enough():
    assert mg.restir_halo_rows() == 20                       # reuse_radius 16 + neighbor_search_radius 4 (ReSTIRDISettings.h:42,65)
    for height, world, halo in ((1080, 8, 20), (96, 2, 20), (64, 8, 20), (40, 3, 20), (1080, 1, 20)):
        tiles = mg.row_tiles(height, world)
        plans = [mg.halo_plan(tiles, r, halo) for r in range(world)]
        for r, plan in enumerate(plans):
            y0, y1 = tiles[r]
            got = set()
            for peer, send, recv in plan:
                # what I send to the peer is what the peer receives from me
                back = [p for p in plans[peer] if p[0] == r]
                assert len(back) == 1 and back[0][2] == send and back[0][1] == recv
                if recv:
                    assert tiles[peer][0] <= recv[0] < recv[1] <= tiles[peer][1]
                    got.update(range(*recv))
            if y1 > y0:
                want = set(range(max(0, y0 - halo), y0)) | set(range(y1, min(height, y1 + halo)))
                assert got == want, (height, world, r)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, height, width, halo, result_queue):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tiles = mg.row_tiles(height, world)
        y0, y1 = tiles[rank]
        rng = np.random.default_rng(7)
        full = rng.integers(0, 256, size=(height, width * 36), dtype=np.uint8)     # 36-byte reservoirs, one row per image row

        # --- halo exchange: start from "only my rows are valid", poison elsewhere ---
        mine = np.full_like(full, 0xAB)
        mine[y0:y1] = full[y0:y1]
        rows = torch.from_numpy(mine)
        mg.exchange_halo(dist, rows, tiles, rank, halo)
        lo, hi = max(0, y0 - halo), min(height, y1 + halo)
        ok_halo = bool(np.array_equal(mine[lo:hi], full[lo:hi]))
        untouched = bool((mine[:lo] == 0xAB).all() and (mine[hi:] == 0xAB).all())

        # --- tile gather: every rank ends with the whole frame ---
        frame = np.zeros((height, width * 12), dtype=np.uint8)
        frame[y0:y1] = full[y0:y1, :width * 12]
        mg.gather_rows(dist, torch.from_numpy(frame), tiles)
        ok_gather = bool(np.array_equal(frame, full[:, :width * 12]))

        # --- column tiles: the same through a strided band ---
        ctiles = mg.column_tiles(width, world)
        x0, x1 = ctiles[rank]
        want_image = rng.random((height, width, 3), dtype=np.float32)
        image = np.zeros_like(want_image)
        image[:, x0:x1] = want_image[:, x0:x1]
        mg.gather_columns(dist, torch.from_numpy(image), ctiles)
        ok_gather = ok_gather and bool(np.array_equal(image, want_image))

        # --- sample-pass sharding: accumulation buffers summed ---
        pixels = torch.full((height, width, 3), float(rank + 1))
        lum = torch.full((height, width), 0.5 * (rank + 1))
        count = torch.full((height, width), rank + 1, dtype=torch.int32)
        total = world * (world + 1) / 2
        ok_reduce = True
        for round_ in range(2):
            # out of place: the rank's own buffers keep their values, so a second reduce (after more rendering) is again the plain sum
            gp, gl, gc = mg.reduce_accumulation(dist, pixels, lum, count)
            scale = round_ + 1
            ok_reduce = ok_reduce and bool((gp == total * scale).all() and (gl == 0.5 * total * scale).all() and (gc == int(total) * scale).all())
            ok_reduce = ok_reduce and bool((pixels == float(rank + 1) * scale).all() and (count == (rank + 1) * scale).all())
            pixels += float(rank + 1); lum += 0.5 * (rank + 1); count += rank + 1          # "another pass rendered"
        result_queue.put((rank, ok_halo, untouched, ok_gather, ok_reduce))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,height", [(2, 96), (3, 64)])
def test_collectives_over_gloo(world, height):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, height, 32, 20, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in results) == list(range(world))
    for rank, ok_halo, untouched, ok_gather, ok_reduce in results:
        assert ok_halo and untouched and ok_gather and ok_reduce, (rank, ok_halo, untouched, ok_gather, ok_reduce)
