#include <vector>
#include <algorithm>
// Multi-GPU plumbing behind the C ABI (include/b200pt_comm.h): tile plan, frame gather, ReSTIR DI halo exchange, accumulation
// all-reduce over NCCL, bound at run time with dlopen so that libb200pt.so has no link-time dependency on libnccl.
// No counterpart in the reference (one device, src/main.cpp:57); SURVEY.md section 8(e).
#include "../../include/b200pt_comm.h"

#include <dlfcn.h>
#include <nccl.h>      // types and prototypes only: the functions are looked up with dlsym

#include <mutex>

namespace
{
struct NcclApi
{
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclBroadcast) Broadcast = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    std::string error;
};

NcclApi& nccl_api()
{
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        // RTLD_NOLOAD first: a copy already in the process (e.g. the one torch bundles) is the one every communicator must share
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names)
            if (!api.handle)
                api.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD);
        for (const char* n : names)
            if (!api.handle)
                api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (!api.handle)
        {
            api.error = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : "");
            return;
        }
#define B200PT_NCCL_SYM(field, name)                                        \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name)); \
    if (!api.field)                                                         \
        api.error += std::string(" missing symbol ") + name;
        B200PT_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        B200PT_NCCL_SYM(CommInitRank, "ncclCommInitRank")
        B200PT_NCCL_SYM(CommDestroy, "ncclCommDestroy")
        B200PT_NCCL_SYM(GetErrorString, "ncclGetErrorString")
        B200PT_NCCL_SYM(Broadcast, "ncclBroadcast")
        B200PT_NCCL_SYM(AllReduce, "ncclAllReduce")
        B200PT_NCCL_SYM(AllGather, "ncclAllGather")
        B200PT_NCCL_SYM(Send, "ncclSend")
        B200PT_NCCL_SYM(Recv, "ncclRecv")
        B200PT_NCCL_SYM(GroupStart, "ncclGroupStart")
        B200PT_NCCL_SYM(GroupEnd, "ncclGroupEnd")
#undef B200PT_NCCL_SYM
    });
    return api;
}

int nccl_fail(ncclResult_t r, const char* what)
{
    NcclApi& api = nccl_api();
    return fail(B200PT_ERROR_CUDA, std::string(what) + ": NCCL error " + (api.GetErrorString ? api.GetErrorString(r) : "?"));
}

#define NCCL_TRY(expr)                       \
    do                                       \
    {                                        \
        ncclResult_t r_ = (expr);            \
        if (r_ != ncclSuccess)               \
            return nccl_fail(r_, #expr);     \
    } while (0)

// [begin, end) of band `rank` of `world` bands over `extent` pixels, multiples of 8 except the last (multigpu.py: row_tiles)
void band_of(int extent, int world, int rank, int& begin, int& end)
{
    const long long blocks = (extent + 7) / 8;
    auto bound = [&](int r) { return r >= world ? extent : (int)std::min<long long>(extent, 8 * ((blocks * r) / world)); };
    begin = bound(rank);
    end = bound(rank + 1);
}
} // namespace

struct b200pt_comm
{
    b200pt_ctx* ctx = nullptr;
    ncclComm_t nccl = nullptr;
    int rank = 0, world = 1;
    unsigned char* staging = nullptr;       // column-band gathers: the bands laid out contiguously, band r at its prefix offset
    size_t staging_bytes = 0;
    double* d_scalars = nullptr;            // b200pt_comm_allreduce_host_f64
    std::vector<int> column_bounds;         // b200pt_comm_set_column_bounds: world + 1 entries, empty = the uniform plan of b200pt_comm_tile

    // column band of rank r of a `width`-pixel image: the bounds set by the caller if they fit this width, else the uniform plan
    void columns_of(int width, int r, int& begin, int& end) const
    {
        if ((int)column_bounds.size() == world + 1 && column_bounds.back() == width)
        {
            begin = column_bounds[r];
            end = column_bounds[r + 1];
        }
        else
            band_of(width, world, r, begin, end);
    }
};

extern "C" {

int b200pt_comm_tile(int sharding, int width, int height, int world, int rank, b200pt_tile* out_tile)
{
    REQUIRE(out_tile != nullptr && width >= 0 && height >= 0 && world >= 1 && rank >= 0 && rank < world);
    REQUIRE(sharding == B200PT_SHARD_ROWS || sharding == B200PT_SHARD_COLUMNS);
    b200pt_tile t = {0, 0, width, height};
    if (sharding == B200PT_SHARD_ROWS)
        band_of(height, world, rank, t.y0, t.y1);
    else
        band_of(width, world, rank, t.x0, t.x1);
    *out_tile = t;
    return B200PT_SUCCESS;
}

int b200pt_comm_balanced_bounds(const int* bounds_in, const double* cost, int world, int extent, int align, int* out_bounds)
{
    REQUIRE(bounds_in != nullptr && cost != nullptr && out_bounds != nullptr && world >= 1 && extent >= 0 && align >= 1);
    const std::vector<int> bounds(bounds_in, bounds_in + world + 1);       // out_bounds may be the same array
    REQUIRE(bounds[0] == 0 && bounds[world] == extent);
    double total = 0.0;
    for (int r = 0; r < world; r++)
    {
        REQUIRE(bounds[r + 1] >= bounds[r] && cost[r] >= 0.0);
        total += bounds[r + 1] > bounds[r] ? cost[r] : 0.0;
    }
    out_bounds[0] = 0;
    out_bounds[world] = extent;
    if (!(total > 0.0))
    {
        for (int r = 1; r < world; r++)
            out_bounds[r] = bounds[r];
        return B200PT_SUCCESS;
    }
    // cost is taken as uniform inside each old band: the cumulative cost is piecewise linear in the pixel coordinate; boundary k goes where it
    // reaches k / world of the total, rounded to a multiple of `align`, kept monotonic and at least one `align` block wide while blocks last
    int band = 0;
    double before = 0.0;       // cumulative cost at bounds[band]
    for (int k = 1; k < world; k++)
    {
        const double target = total * (double)k / (double)world;
        while (band < world - 1 && (bounds[band + 1] == bounds[band] || before + cost[band] < target))
        {
            before += bounds[band + 1] > bounds[band] ? cost[band] : 0.0;
            band++;
        }
        const double width = (double)(bounds[band + 1] - bounds[band]);
        const double x = cost[band] > 0.0 ? bounds[band] + width * std::min(1.0, std::max(0.0, (target - before) / cost[band])) : (double)bounds[band];
        long long rounded = (long long)((x + 0.5 * align) / align) * align;
        const long long lowest = std::min<long long>(extent, (long long)out_bounds[k - 1] + align);
        const long long highest = std::max<long long>(lowest, (long long)extent - (long long)(world - k) * align);
        rounded = std::min(std::max(rounded, lowest), highest);
        out_bounds[k] = (int)std::min<long long>(rounded, extent);
    }
    return B200PT_SUCCESS;
}

int b200pt_comm_set_column_bounds(b200pt_comm* comm, const int* bounds)
{
    REQUIRE(comm != nullptr);
    if (bounds == nullptr)
    {
        comm->column_bounds.clear();
        return B200PT_SUCCESS;
    }
    REQUIRE(bounds[0] == 0);
    for (int r = 0; r < comm->world; r++)
        REQUIRE(bounds[r + 1] >= bounds[r]);
    comm->column_bounds.assign(bounds, bounds + comm->world + 1);
    return B200PT_SUCCESS;
}

int b200pt_comm_restir_halo_rows(int reuse_radius, int neighbor_search_radius, int camera_motion_rows)
{
    return (reuse_radius > 0 ? reuse_radius : 0) + (neighbor_search_radius > 0 ? neighbor_search_radius : 0) + (camera_motion_rows > 0 ? camera_motion_rows : 0);
}

int b200pt_comm_get_unique_id(void* out_id_128_bytes)
{
    REQUIRE(out_id_128_bytes != nullptr);
    NcclApi& api = nccl_api();
    if (!api.error.empty())
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "NCCL is not available: " + api.error);
    static_assert(sizeof(ncclUniqueId) == B200PT_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NCCL_TRY(api.GetUniqueId(&id));
    memcpy(out_id_128_bytes, &id, sizeof(id));
    return B200PT_SUCCESS;
}

int b200pt_comm_create(b200pt_ctx* ctx, int rank, int world, const void* id_128_bytes, b200pt_comm** out_comm)
{
    REQUIRE(ctx != nullptr && out_comm != nullptr && id_128_bytes != nullptr && world >= 1 && rank >= 0 && rank < world);
    *out_comm = nullptr;
    NcclApi& api = nccl_api();
    if (!api.error.empty())
        return fail(B200PT_ERROR_INVALID_ARGUMENT, "NCCL is not available: " + api.error);
    CUDA_TRY(cudaSetDevice(ctx->device));
    ncclUniqueId id;
    memcpy(&id, id_128_bytes, sizeof(id));
    b200pt_comm* comm = new b200pt_comm();
    comm->ctx = ctx;
    comm->rank = rank;
    comm->world = world;
    ncclResult_t r = api.CommInitRank(&comm->nccl, world, id, rank);
    if (r != ncclSuccess)
    {
        delete comm;
        return nccl_fail(r, "ncclCommInitRank");
    }
    cudaError_t e = cudaMalloc(&comm->d_scalars, sizeof(double) * 64);
    if (e != cudaSuccess)
    {
        api.CommDestroy(comm->nccl);
        delete comm;
        return cuda_fail(e, "cudaMalloc(comm scalars)");
    }
    *out_comm = comm;
    return B200PT_SUCCESS;
}

int b200pt_comm_destroy(b200pt_comm* comm)
{
    if (!comm)
        return B200PT_SUCCESS;
    cudaSetDevice(comm->ctx->device);
    cudaStreamSynchronize(comm->ctx->main_stream);
    nccl_api().CommDestroy(comm->nccl);
    cudaFree(comm->staging);
    cudaFree(comm->d_scalars);
    delete comm;
    return B200PT_SUCCESS;
}

int b200pt_comm_rank(const b200pt_comm* comm) { return comm ? comm->rank : -1; }
int b200pt_comm_world(const b200pt_comm* comm) { return comm ? comm->world : 0; }

} // extern "C"

extern "C" {

int b200pt_comm_gather_tiles(b200pt_comm* comm, void* d_image, int width, int height, int bytes_per_pixel, int sharding, b200pt_stream stream)
{
    REQUIRE(comm != nullptr && d_image != nullptr && width > 0 && height > 0 && bytes_per_pixel > 0);
    REQUIRE(sharding == B200PT_SHARD_ROWS || sharding == B200PT_SHARD_COLUMNS);
    NcclApi& api = nccl_api();
    cudaStream_t st = stream_of(comm->ctx, stream);
    unsigned char* image = static_cast<unsigned char*>(d_image);
    const size_t row_bytes = (size_t)width * bytes_per_pixel;
    if (sharding == B200PT_SHARD_ROWS)
    {
        // bands are contiguous: every band is broadcast in place by its owner, as one grouped operation
        NCCL_TRY(api.GroupStart());
        for (int r = 0; r < comm->world; r++)
        {
            int y0, y1;
            band_of(height, comm->world, r, y0, y1);
            if (y1 > y0)
            {
                unsigned char* band = image + (size_t)y0 * row_bytes;
                ncclResult_t res = api.Broadcast(band, band, (size_t)(y1 - y0) * row_bytes, ncclUint8, r, comm->nccl, st);
                if (res != ncclSuccess)
                {
                    api.GroupEnd();
                    return nccl_fail(res, "ncclBroadcast(row band)");
                }
            }
        }
        NCCL_TRY(api.GroupEnd());
        return B200PT_SUCCESS;
    }
    // column bands are strided: pack mine into the staging buffer (band r at the prefix offset of its bytes), broadcast every band
    // from its owner, unpack the others'
    const size_t total = row_bytes * (size_t)height;
    if (comm->staging_bytes < total)
    {
        CUDA_TRY(cudaStreamSynchronize(st));
        cudaFree(comm->staging);
        comm->staging = nullptr;
        comm->staging_bytes = 0;
        CUDA_TRY(cudaMalloc(&comm->staging, total));
        comm->staging_bytes = total;
    }
    int x0, x1;
    comm->columns_of(width, comm->rank, x0, x1);
    if (x1 > x0)
        CUDA_TRY(cudaMemcpy2DAsync(comm->staging + (size_t)x0 * bytes_per_pixel * height, (size_t)(x1 - x0) * bytes_per_pixel, image + (size_t)x0 * bytes_per_pixel, row_bytes,
                                   (size_t)(x1 - x0) * bytes_per_pixel, (size_t)height, cudaMemcpyDeviceToDevice, st));
    NCCL_TRY(api.GroupStart());
    for (int r = 0; r < comm->world; r++)
    {
        int a, b;
        comm->columns_of(width, r, a, b);
        if (b > a)
        {
            unsigned char* band = comm->staging + (size_t)a * bytes_per_pixel * height;
            ncclResult_t res = api.Broadcast(band, band, (size_t)(b - a) * bytes_per_pixel * height, ncclUint8, r, comm->nccl, st);
            if (res != ncclSuccess)
            {
                api.GroupEnd();
                return nccl_fail(res, "ncclBroadcast(column band)");
            }
        }
    }
    NCCL_TRY(api.GroupEnd());
    for (int r = 0; r < comm->world; r++)
    {
        int a, b;
        comm->columns_of(width, r, a, b);
        if (b > a && r != comm->rank)
            CUDA_TRY(cudaMemcpy2DAsync(image + (size_t)a * bytes_per_pixel, row_bytes, comm->staging + (size_t)a * bytes_per_pixel * height, (size_t)(b - a) * bytes_per_pixel,
                                       (size_t)(b - a) * bytes_per_pixel, (size_t)height, cudaMemcpyDeviceToDevice, st));
    }
    return B200PT_SUCCESS;
}

int b200pt_comm_exchange_halo(b200pt_comm* comm, void* d_rows, int height, size_t row_bytes, int halo_rows, b200pt_stream stream)
{
    REQUIRE(comm != nullptr && d_rows != nullptr && height > 0 && row_bytes > 0 && halo_rows >= 0);
    NcclApi& api = nccl_api();
    cudaStream_t st = stream_of(comm->ctx, stream);
    unsigned char* rows = static_cast<unsigned char*>(d_rows);
    int y0, y1;
    band_of(height, comm->world, comm->rank, y0, y1);
    if (y1 <= y0 || halo_rows == 0 || comm->world == 1)
        return B200PT_SUCCESS;
    const int lo = std::max(0, y0 - halo_rows), hi = std::min(height, y1 + halo_rows);
    NCCL_TRY(api.GroupStart());
    for (int peer = 0; peer < comm->world; peer++)
    {
        int p0, p1;
        band_of(height, comm->world, peer, p0, p1);
        if (peer == comm->rank || p1 <= p0)
            continue;
        // rows of the peer inside my halo: I receive them; rows of mine inside the peer's halo: I send them (multigpu.py: halo_plan)
        const int r0 = std::max(p0, lo), r1 = std::min(p1, hi);
        const int s0 = std::max(y0, p0 - halo_rows), s1 = std::min(y1, p1 + halo_rows);
        ncclResult_t res = ncclSuccess;
        if (s1 > s0)
            res = api.Send(rows + (size_t)s0 * row_bytes, (size_t)(s1 - s0) * row_bytes, ncclUint8, peer, comm->nccl, st);
        if (res == ncclSuccess && r1 > r0)
            res = api.Recv(rows + (size_t)r0 * row_bytes, (size_t)(r1 - r0) * row_bytes, ncclUint8, peer, comm->nccl, st);
        if (res != ncclSuccess)
        {
            api.GroupEnd();
            return nccl_fail(res, "ncclSend/ncclRecv(halo rows)");
        }
    }
    NCCL_TRY(api.GroupEnd());
    return B200PT_SUCCESS;
}

int b200pt_comm_allreduce_sum(b200pt_comm* comm, const void* d_in, void* d_out, size_t count, int dtype, b200pt_stream stream)
{
    REQUIRE(comm != nullptr && (count == 0 || (d_in != nullptr && d_out != nullptr)) && (dtype == 0 || dtype == 1));
    if (count == 0)
        return B200PT_SUCCESS;
    NCCL_TRY(nccl_api().AllReduce(d_in, d_out, count, dtype == 0 ? ncclFloat32 : ncclInt32, ncclSum, comm->nccl, stream_of(comm->ctx, stream)));
    return B200PT_SUCCESS;
}

int b200pt_comm_allreduce_host_f64(b200pt_comm* comm, double* inout, int count, int op)
{
    REQUIRE(comm != nullptr && inout != nullptr && count >= 1 && count <= 64 && (op == 0 || op == 1));
    cudaStream_t st = comm->ctx->main_stream;
    CUDA_TRY(cudaMemcpyAsync(comm->d_scalars, inout, sizeof(double) * count, cudaMemcpyHostToDevice, st));
    NCCL_TRY(nccl_api().AllReduce(comm->d_scalars, comm->d_scalars, (size_t)count, ncclFloat64, op == 0 ? ncclSum : ncclMax, comm->nccl, st));
    CUDA_TRY(cudaMemcpyAsync(inout, comm->d_scalars, sizeof(double) * count, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return B200PT_SUCCESS;
}

int b200pt_comm_barrier(b200pt_comm* comm)
{
    double zero = 0.0;
    return b200pt_comm_allreduce_host_f64(comm, &zero, 1, 0);
}

} // extern "C"
