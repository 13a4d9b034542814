// One FullPathTracer permutation per translation unit: compile with -DPT_LSS= -DPT_ESS= -DPT_LATER= -DPT_GGX=
// Holds the megakernel and the wavefront kernels (wavefront.cuh) of the permutation plus their launcher.
#include "wavefront.cuh"

namespace pt
{
__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_FullPathTracer)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    full_path_tracer_pixel<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, x, y);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_bounce0)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    wf_bounce0_pixel<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, tile, wf, x, y);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_bounce)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wf.counts[2 * bounce])
        return;
    wf_bounce_path<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, tile, wf, bounce, i);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_gen)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf)
{
    const uint32_t x = tile.x0 + blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = tile.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= (uint32_t)tile.x1 || y >= (uint32_t)tile.y1)
        return;
    wf_gen_pixel<PT_ESS>(rd, res, tile, wf, x, y);
}

__global__ void __launch_bounds__(128) PT_VARIANT_NAME(k_wf_extend)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wf.counts[2 * bounce])
        return;
    wf_extend_path<PT_ESS>(rd, res, tile, wf, bounce, i);
}

__global__ void __launch_bounds__(kSortBuckets) PT_VARIANT_NAME(k_wf_scan)(const __grid_constant__ WavefrontState wf)
{
    __shared__ uint32_t smem[kSortBuckets];
    wf_scan_buckets(wf, smem);
}

__global__ void __launch_bounds__(256) PT_VARIANT_NAME(k_wf_scatter)(const __grid_constant__ WavefrontState wf, int bounce)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wf.counts[2 * bounce + 1])
        return;
    wf_scatter_path(wf, i);
}

__global__ void __launch_bounds__(64) PT_VARIANT_NAME(k_wf_shade)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wf.counts[2 * bounce + 1])
        return;
    wf_shade_path<PT_LSS, PT_ESS, PT_LATER, PT_GGX>(rd, res, tile, wf, bounce, i);
}

cudaError_t PT_VARIANT_NAME(launch_FullPathTracer)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream, const WavefrontState* wf, int mode)
{
    dim3 grid((tile.x1 - tile.x0 + 7) / 8, (tile.y1 - tile.y0 + 7) / 8, 1);
    if (mode == WF_MEGAKERNEL || wf == nullptr)
    {
        PT_VARIANT_NAME(k_FullPathTracer)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
        return cudaGetLastError();
    }

    int nb_bounces = rd->render_settings.nb_bounces;
    if (rd->render_settings.wants_render_low_resolution && rd->render_settings.allow_render_low_resolution && rd->render_settings.accumulate)
        nb_bounces = nb_bounces < 3 ? nb_bounces : 3;
    const uint32_t paths = (uint32_t)(tile.x1 - tile.x0) * (uint32_t)(tile.y1 - tile.y0);
    if (paths > wf->capacity || (uint32_t)nb_bounces > wf->max_bounces)
        return cudaErrorInvalidValue;

    cudaError_t e = cudaMemsetAsync(wf->counts, 0, sizeof(uint32_t) * 2 * (wf->max_bounces + 2), stream);
    if (e != cudaSuccess)
        return e;
    if (mode == WF_FUSED)
    {
        PT_VARIANT_NAME(k_wf_bounce0)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *wf);
        for (int bounce = 1; bounce <= nb_bounces; bounce++)
            PT_VARIANT_NAME(k_wf_bounce)<<<(paths + 63) / 64, 64, 0, stream>>>(*rd, res, tile, *wf, bounce);
        return cudaGetLastError();
    }
    for (int bounce = 0; bounce <= nb_bounces; bounce++)
    {
        if (bounce == 0)
            PT_VARIANT_NAME(k_wf_gen)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile, *wf);
        else
            PT_VARIANT_NAME(k_wf_extend)<<<(paths + 127) / 128, 128, 0, stream>>>(*rd, res, tile, *wf, bounce);
        PT_VARIANT_NAME(k_wf_scan)<<<1, kSortBuckets, 0, stream>>>(*wf);
        PT_VARIANT_NAME(k_wf_scatter)<<<(paths + 255) / 256, 256, 0, stream>>>(*wf, bounce);
        // the scatter cursors become the next histogram: zero them again
        e = cudaMemsetAsync(wf->bucket_count, 0, sizeof(uint32_t) * kSortBuckets, stream);
        if (e != cudaSuccess)
            return e;
        PT_VARIANT_NAME(k_wf_shade)<<<(paths + 63) / 64, 64, 0, stream>>>(*rd, res, tile, *wf, bounce);
    }
    return cudaGetLastError();
}

} // namespace pt
