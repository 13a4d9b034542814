/*
 * b200pt_comm.h -- multi-GPU plumbing of the render hot path behind the C ABI (part of libb200pt.so).
 *
 * The reference renders on ONE device (src/main.cpp:57; HIPRTOrochiCtx::init(0), src/HIPRT-Orochi/HIPRTOrochiCtx.h:29-57) and has no
 * distributed backend: nothing here replaces a reference call site. It is what SURVEY.md section 8(e) adds around the launch points of
 * GPURenderer::render (src/Renderer/GPURenderer.cpp:408-486): one process per GPU, scene + BVH8 replicated, the frame sharded by image
 * tiles (bit-identical to the one-GPU render: pixel indices and seeds stay global) or by sample passes, and
 *   - the frame gathered once per OUTPUT (not per sample)                          b200pt_comm_gather_tiles
 *   - ReSTIR DI output reservoirs of the halo rows exchanged once per frame        b200pt_comm_exchange_halo
 *   - accumulation buffers summed over ranks (sample-pass sharding)                b200pt_comm_allreduce_sum
 * over NCCL (NVLink 5 / NVSwitch on the 8xB200 box). NCCL is bound at run time (dlopen of libnccl.so.2, the copy already loaded in the
 * process if there is one), so libb200pt.so itself has no link-time dependency on it; b200pt_comm_create fails with
 * B200PT_ERROR_INVALID_ARGUMENT and a message if the library cannot be found. Rendezvous is the launcher's business: rank 0 calls
 * b200pt_comm_get_unique_id and hands the 128 bytes to the other ranks by whatever channel it has (a file, MPI, torch.distributed).
 * All collectives are asynchronous on the given stream (NULL = the context's main stream), ordered with the kernel launches on it.
 */
#ifndef B200PT_COMM_H
#define B200PT_COMM_H

#include "b200pt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200pt_comm b200pt_comm;
#define B200PT_COMM_ID_BYTES 128

typedef enum b200pt_sharding
{
    B200PT_SHARD_ROWS = 0,       /* contiguous row bands (multiples of 8 rows): ReSTIR DI (halo rows), contiguous in memory */
    B200PT_SHARD_COLUMNS = 1     /* contiguous column bands (multiples of 8 columns): balanced on landscape frames (sky above, geometry below) */
} b200pt_sharding;

/* The band of `rank` among `world` bands of a width x height image; bands tile the image, are multiples of 8 pixels except the last,
 * and may be empty when world exceeds the number of 8-pixel blocks. No communicator needed. */
B200PT_API int b200pt_comm_tile(int sharding, int width, int height, int world, int rank, b200pt_tile* out_tile);
/* Cost-balanced column bands (new). A landscape frame's columns do not cost the same (path lengths differ): after a frame every rank
 * reports what its band cost (e.g. the milliseconds of its launches, shared with b200pt_comm_allreduce_host_f64) and all ranks compute the same
 * new plan. bounds / out_bounds: world + 1 entries, band r = [bounds[r], bounds[r + 1]), bounds[0] = 0, bounds[world] = extent. The cost is
 * taken as uniform inside each old band; new boundaries are multiples of `align` pixels, bands stay at least `align` wide while the image has
 * that many blocks; out_bounds may be the bounds array itself. Iterate over a few frames. No communicator needed. Tiles only change which rank renders a pixel, never its value. */
B200PT_API int b200pt_comm_balanced_bounds(const int* bounds, const double* cost, int world, int extent, int align, int* out_bounds);
/* The column plan b200pt_comm_gather_tiles uses from now on for images `bounds[world]` pixels wide (NULL: back to b200pt_comm_tile's). */
B200PT_API int b200pt_comm_set_column_bounds(b200pt_comm* comm, const int* bounds);
/* Rows of neighbour data the fused ReSTIR DI pass can reach above and below a tile: reuse_radius + neighbor_search_radius
 * (ReSTIRDISettings.h:42,65; Utils.h:301-366,426-478) + the camera's motion in rows since the last frame. */
B200PT_API int b200pt_comm_restir_halo_rows(int reuse_radius, int neighbor_search_radius, int camera_motion_rows);

B200PT_API int b200pt_comm_get_unique_id(void* out_id_128_bytes);
B200PT_API int b200pt_comm_create(b200pt_ctx* ctx, int rank, int world, const void* id_128_bytes, b200pt_comm** out_comm);
B200PT_API int b200pt_comm_destroy(b200pt_comm* comm);
B200PT_API int b200pt_comm_rank(const b200pt_comm* comm);
B200PT_API int b200pt_comm_world(const b200pt_comm* comm);

/* d_image: width x height pixels of bytes_per_pixel bytes (e.g. 12 for RenderBuffers::pixels), row-major; this rank has written the band
 * b200pt_comm_tile gives it. On completion every rank holds the whole image. Row bands travel in place; column bands (strided) go
 * through a staging buffer the communicator owns (grown on demand, kept). */
B200PT_API int b200pt_comm_gather_tiles(b200pt_comm* comm, void* d_image, int width, int height, int bytes_per_pixel, int sharding, b200pt_stream stream);
/* d_rows: height rows of row_bytes bytes (e.g. width * 36 for ReSTIRDIReservoir), row band sharding; this rank has written its own rows.
 * On completion the rows within `halo_rows` of its band hold their owners' values (a halo wider than a neighbour's band reaches the
 * next rank too). One grouped batch of ncclSend / ncclRecv. */
B200PT_API int b200pt_comm_exchange_halo(b200pt_comm* comm, void* d_rows, int height, size_t row_bytes, int halo_rows, b200pt_stream stream);
/* d_out[i] = sum over ranks of d_in[i]; d_out may be d_in (in place: every rank then holds the global sum and must not add it again)
 * or a separate buffer (the rank's own accumulation stays valid for further rendering). dtype: 0 = float32, 1 = int32. */
B200PT_API int b200pt_comm_allreduce_sum(b200pt_comm* comm, const void* d_in, void* d_out, size_t count, int dtype, b200pt_stream stream);
/* Host scalars through the communicator (timing: max over ranks of a device-timed interval; totals: sum). op: 0 = sum, 1 = max.
 * Synchronises the stream. */
B200PT_API int b200pt_comm_allreduce_host_f64(b200pt_comm* comm, double* inout, int count, int op);
B200PT_API int b200pt_comm_barrier(b200pt_comm* comm);

#ifdef __cplusplus
}
#endif

#endif /* B200PT_COMM_H */
