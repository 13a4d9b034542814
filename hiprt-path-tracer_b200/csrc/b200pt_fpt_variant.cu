// One FullPathTracer permutation per translation unit: compile with -DPT_LSS= -DPT_ESS= -DPT_LATER= -DPT_GGX=
#include "kernels.cuh"
#include "launch.h"

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

cudaError_t PT_VARIANT_NAME(launch_FullPathTracer)(const HIPRTRenderData* rd, int2 res, Tile tile, cudaStream_t stream)
{
    dim3 grid((tile.x1 - tile.x0 + 7) / 8, (tile.y1 - tile.y0 + 7) / 8, 1);
    PT_VARIANT_NAME(k_FullPathTracer)<<<grid, dim3(8, 8, 1), 0, stream>>>(*rd, res, tile);
    return cudaGetLastError();
}

} // namespace pt
