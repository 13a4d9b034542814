"""Multi-GPU plumbing of the render path: one process per GPU, scene + BVH8 replicated, the image sharded by
row tiles or by sample passes. The reference has no distributed backend (single device, src/main.cpp:57); this is new.

  * tile sharding      rank r renders rows [y0_r, y1_r) (or columns [x0_r, x1_r): `column_tiles`, the balanced choice for landscape
                       frames) with the global pixel_index = x + y*W, so its pixels are bit-identical to the 1-GPU render;
                       `gather_rows` / `gather_columns` assemble the frame once per output (not per sample).
  * sample-pass        every rank renders the full frame for disjoint sample numbers; `reduce_accumulation` sums the accumulation
    sharding           buffers (RenderData.h:34-36 pixels, AuxiliaryBuffers pixel_squared_luminance / pixel_sample_count).
  * ReSTIR DI halo     the fused spatiotemporal pass reads last frame's output reservoirs within reuse_radius of the temporal
                       position, itself within neighbor_search_radius of the pixel (ReSTIRDISettings.h:42,65; Utils.h:301-366,
                       426-478): `restir_halo_rows` rows either side of a tile. The G-buffer and this frame's initial candidates
                       are deterministic per pixel, so each rank recomputes CameraRays over tile + halo itself
                       (GPURenderer::set_tile_halo) and only the 36-byte reservoirs of the halo rows travel (`exchange_halo`).

The collectives are torch.distributed calls on tensors that alias the renderer's device buffers (`device_rows`): NCCL over NVLink on
the GPUs, gloo on CPU tensors in tests/test_multigpu_gloo.py. Buffers are viewed as [H, row_bytes] so that a band of rows is one
contiguous slice.
"""
import torch


def row_tiles(height, world, align=8):
    """Contiguous row bands, one per rank, each a multiple of `align` rows (the 8x8 launch blocks) except the last."""
    if world < 1 or height < 0:
        raise ValueError("row_tiles: world must be >= 1 and height >= 0")
    blocks = (height + align - 1) // align
    bounds = [min(height, align * ((blocks * r) // world)) for r in range(world)] + [height]
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def column_tiles(width, world, align=8):
    """Contiguous column bands, one per rank, each a multiple of `align` pixels except the last. A landscape frame is sky above
    and geometry below, so row bands give the ranks very different work (C3 on 2 GPUs: 12.0 vs 17.4 ms per pass); column bands
    cut across that gradient (the path lanes of one GPU are column bands for the same reason, GPURenderer::launch_path_tracing)."""
    return row_tiles(width, world, align)


def gather_columns(dist, image, tiles):
    """image: [H, W, C] tensor aliasing a per-pixel buffer whose columns tiles[rank] this rank has written. Every rank ends with
    the whole frame: each band is broadcast by its owner through a contiguous staging copy (a column band is strided in memory)."""
    for owner, (x0, x1) in enumerate(tiles):
        if x1 > x0:
            band = image[:, x0:x1].contiguous()
            dist.broadcast(band, src=owner)
            image[:, x0:x1] = band


def restir_halo_rows(reuse_radius=16, neighbor_search_radius=4, camera_motion_rows=0):
    """Rows of neighbour data the fused ReSTIR DI pass can reach above and below a tile."""
    return int(reuse_radius) + int(neighbor_search_radius) + int(camera_motion_rows)


def halo_plan(tiles, rank, halo):
    """[(peer, (send_y0, send_y1), (recv_y0, recv_y1))]: what `rank` sends from its own rows and receives into its halo rows.
    A halo wider than a neighbour's tile reaches the next rank too, so every rank whose rows fall inside the halo is a peer."""
    y0, y1 = tiles[rank]
    if y1 <= y0 or halo <= 0:
        return []
    height = tiles[-1][1]
    lo, hi = max(0, y0 - halo), min(height, y1 + halo)
    plan = []
    for peer, (p0, p1) in enumerate(tiles):
        if peer == rank or p1 <= p0:
            continue
        # rows of the peer inside my halo: I receive them
        r0, r1 = max(p0, lo), min(p1, hi)
        # rows of mine inside the peer's halo: I send them
        s0, s1 = max(y0, p0 - halo), min(y1, p1 + halo)
        recv = (r0, r1) if r1 > r0 else None
        send = (s0, s1) if s1 > s0 else None
        if recv or send:
            plan.append((peer, send, recv))
    return plan


def exchange_halo(dist, rows, tiles, rank, halo):
    """rows: [H, row_bytes] tensor aliasing a per-pixel buffer whose rows tiles[rank] this rank has just written. After the call
    the rows within `halo` of the tile hold the owners' values. One grouped batch of point-to-point operations."""
    ops = []
    for peer, send, recv in halo_plan(tiles, rank, halo):
        if send:
            ops.append(dist.P2POp(dist.isend, rows[send[0]:send[1]], peer))
        if recv:
            ops.append(dist.P2POp(dist.irecv, rows[recv[0]:recv[1]], peer))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return len(ops)


def gather_rows(dist, rows, tiles):
    """Every rank ends with the whole [H, row_bytes] buffer. Equal bands: ONE all_gather (in place: the output list aliases the bands
    of `rows`); ragged bands (the last one is shorter when H is not a multiple of 8 * world): one broadcast per band."""
    sizes = {y1 - y0 for y0, y1 in tiles}
    if len(sizes) == 1 and next(iter(sizes)) > 0:
        rank = dist.get_rank()
        y0, y1 = tiles[rank]
        dist.all_gather([rows[a:b] for a, b in tiles], rows[y0:y1].clone())
        return
    for owner, (y0, y1) in enumerate(tiles):
        if y1 > y0:
            dist.broadcast(rows[y0:y1], src=owner)


def reduce_accumulation(dist, pixels, squared_luminance=None, sample_count=None):
    """Sample-pass sharding: the sum over ranks of the accumulation buffers (RenderData.h:34-36 `pixels`, and optionally
    pixel_squared_luminance / pixel_sample_count), returned as NEW tensors. The ranks' own buffers are left untouched, so the call can be
    repeated at any time (every K passes, or once per output) and rendering can go on afterwards: an in-place reduce would make every
    rank hold the global sum and count it again at the next reduce. Consequence, stated plainly: each rank's adaptive-sampling
    decisions (AdaptiveSampling.h:30-104) are taken on ITS OWN statistics, not on the global ones; the reduced image divided by the
    reduced sample count is the global estimate. Float summation order changes: statistically, not bitwise, equal to one GPU."""
    out = []
    for t in (pixels, squared_luminance, sample_count):
        if t is None:
            out.append(None)
            continue
        total = t.clone()
        dist.all_reduce(total, op=dist.ReduceOp.SUM)
        out.append(total)
    return tuple(out)


class _DevicePointer:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (int(ptr), False), "version": 3}


def device_rows(ptr, height, row_bytes, device):
    """[H, row_bytes] uint8 tensor aliasing a renderer device buffer (no copy)."""
    if not ptr:
        raise ValueError("device_rows: null device pointer")
    return torch.as_tensor(_DevicePointer(ptr, (int(height), int(row_bytes)), "|u1"), device=device)


def device_array(ptr, shape, typestr, device):
    """Typed tensor aliasing a renderer device buffer, e.g. the float32 [H, W, 3] accumulation buffer."""
    if not ptr:
        raise ValueError("device_array: null device pointer")
    return torch.as_tensor(_DevicePointer(ptr, tuple(int(s) for s in shape), typestr), device=device)


class Comm:
    """b200pt_comm (include/b200pt_comm.h): the same plumbing behind the C ABI, over NCCL bound by libb200pt.so itself -- what a C/C++
    host links against. Python is only the launcher here: `unique_id` is made on rank 0 and handed to the other ranks by the caller
    (bench.py broadcasts it with torch.distributed, a C++ host would use a file or MPI)."""
    ROWS, COLUMNS = 0, 1

    def __init__(self, ctx, rank, world, unique_id):
        import ctypes as C
        from . import _capi
        self._C, self._capi, self.ctx, self.lib = C, _capi, ctx, ctx.lib
        self.handle = C.c_void_p()
        ident = (C.c_ubyte * 128).from_buffer_copy(bytes(unique_id))
        _capi.check(self.lib.b200pt_comm_create(ctx.handle, C.c_int(rank), C.c_int(world), ident, C.byref(self.handle)), "b200pt_comm_create")
        self.rank, self.world = rank, world

    @staticmethod
    def unique_id(lib):
        import ctypes as C
        from . import _capi
        ident = (C.c_ubyte * 128)()
        _capi.check(lib.b200pt_comm_get_unique_id(ident), "b200pt_comm_get_unique_id")
        return bytes(ident)

    @staticmethod
    def tile(lib, sharding, width, height, world, rank):
        """(x0, y0, x1, y1) of the rank's band (b200pt_comm_tile)."""
        import ctypes as C
        from . import _capi, structs
        t = structs.Tile()
        _capi.check(lib.b200pt_comm_tile(C.c_int(sharding), C.c_int(width), C.c_int(height), C.c_int(world), C.c_int(rank), C.byref(t)), "b200pt_comm_tile")
        return (t.x0, t.y0, t.x1, t.y1)

    def close(self):
        if self.handle:
            self.lib.b200pt_comm_destroy(self.handle)
            self.handle = self._C.c_void_p()

    def gather_tiles(self, dptr, width, height, bytes_per_pixel, sharding):
        C = self._C
        self._capi.check(self.lib.b200pt_comm_gather_tiles(self.handle, C.c_void_p(dptr), C.c_int(width), C.c_int(height), C.c_int(bytes_per_pixel), C.c_int(sharding), None), "b200pt_comm_gather_tiles")

    def exchange_halo(self, dptr, height, row_bytes, halo):
        C = self._C
        self._capi.check(self.lib.b200pt_comm_exchange_halo(self.handle, C.c_void_p(dptr), C.c_int(height), C.c_size_t(row_bytes), C.c_int(halo), None), "b200pt_comm_exchange_halo")

    def allreduce_sum(self, d_in, d_out, count, dtype="f32"):
        C = self._C
        self._capi.check(self.lib.b200pt_comm_allreduce_sum(self.handle, C.c_void_p(d_in), C.c_void_p(d_out), C.c_size_t(count), C.c_int(0 if dtype == "f32" else 1), None), "b200pt_comm_allreduce_sum")

    def allreduce_host(self, values, op="sum"):
        C = self._C
        arr = (C.c_double * len(values))(*[float(v) for v in values])
        self._capi.check(self.lib.b200pt_comm_allreduce_host_f64(self.handle, arr, C.c_int(len(values)), C.c_int(0 if op == "sum" else 1)), "b200pt_comm_allreduce_host_f64")
        return list(arr)

    def barrier(self):
        self._capi.check(self.lib.b200pt_comm_barrier(self.handle), "b200pt_comm_barrier")

    @staticmethod
    def balanced_bounds(lib, bounds, costs, extent, align=8):
        """b200pt_comm_balanced_bounds: new band boundaries at which the bands would cost the same (cost uniform inside each old band)."""
        import ctypes as C
        from . import _capi
        world = len(costs)
        assert len(bounds) == world + 1
        out = (C.c_int * (world + 1))()
        _capi.check(lib.b200pt_comm_balanced_bounds((C.c_int * (world + 1))(*[int(b) for b in bounds]), (C.c_double * world)(*[float(c) for c in costs]), C.c_int(world),
                                                    C.c_int(extent), C.c_int(align), out), "b200pt_comm_balanced_bounds")
        return list(out)

    def set_column_bounds(self, bounds):
        """The column plan gather_tiles uses from now on (None: the uniform plan)."""
        C = self._C
        arr = None if bounds is None else (C.c_int * len(bounds))(*[int(b) for b in bounds])
        self._capi.check(self.lib.b200pt_comm_set_column_bounds(self.handle, arr), "b200pt_comm_set_column_bounds")

    def rebalance_columns(self, bounds, my_cost, extent, align=8):
        """One balancing step: every rank contributes the cost of its band, all compute the same new plan and adopt it."""
        costs = [0.0] * self.world
        costs[self.rank] = float(my_cost)
        costs = self.allreduce_host(costs)
        new = Comm.balanced_bounds(self.lib, bounds, costs, extent, align)
        self.set_column_bounds(new)
        return new, costs

