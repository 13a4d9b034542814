// The phase kernels P1..P5 of the phased MIS pipeline (wavefront_mis.cuh) of one LSS_MIS_LIGHT_BSDF permutation, in a translation
// unit of their own: it is compiled with -DPT_BSDF_INLINE, so that principled_bsdf_eval / _sample are inlined into the one place
// each phase calls them and the compiler keeps the material fields a phase reads in registers instead of a 256-byte local copy
// (stack frames 700 -> 240-270 bytes, shade stage -13 %; profiles/r01_experiments_shade_and_trace.txt). The other kernels of
// the permutation keep the BSDF out of line (b200pt_fpt_variant.cu): inlined at their many call sites they take 9x longer to build.
#include "wavefront_mis.cuh"

namespace pt
{
#define PT_FOR_EACH_QUEUED(I, COUNT, KID)                                                                           \
    const uint32_t n_queued_ = (COUNT);                                                                             \
    uint32_t* const cursor_ = wf.cursors + (uint32_t)bounce * kCursorsPerBounce + (KID);                            \
    for (uint32_t I = wf_next_entry(cursor_); __any_sync(0xFFFFFFFFu, I < n_queued_); I = wf_next_entry(cursor_))   \
        if (I < n_queued_)

// The phase kernels wait on memory (ncu, profiles/r01_shade_phases_c3.txt: issue-active 18-23 %, long-scoreboard 7-12 stalled warps
// per issue at 114-118 registers = 16 warps per SM): cap the registers so that more warps are resident to cover the latency.
#ifndef PT_PHASE_BLOCKS_PER_SM
#define PT_PHASE_BLOCKS_PER_SM 12
#endif
#define PT_QUEUE_KERNEL(NAME, BOUNDS, KID, CALL)                                                                                                                            \
    __global__ void __launch_bounds__(BOUNDS, PT_PHASE_BLOCKS_PER_SM) PT_VARIANT_NAME(NAME)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce) \
    {                                                                                                                                                                  \
        PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce + 1], KID)                                                                                                          \
            CALL(rd, res, tile, wf, bounce, i);                                                                                                                        \
    }
#if PT_LSS == 3
PT_QUEUE_KERNEL(k_wf_nee_light_eval, 64, 2, nee_light_eval_path)
PT_QUEUE_KERNEL(k_wf_nee_bsdf_sample, 64, 3, nee_bsdf_sample_path<PT_GGX>)
#endif
PT_QUEUE_KERNEL(k_wf_env_light_eval, 64, 4, env_light_eval_path<PT_ESS>)
PT_QUEUE_KERNEL(k_wf_env_bsdf_sample, 64, 5, env_bsdf_sample_path<PT_GGX>)
// eval + sample of a MIS pair on one loaded vertex (B200PT_PAIR_PHASES=1); they take the queue cursor of the eval phase
#define PT_NEE_PAIR nee_pair_path<PT_GGX>
#define PT_ENV_PAIR env_pair_path<PT_ESS, PT_GGX>
#if PT_LSS == 3
PT_QUEUE_KERNEL(k_wf_nee_pair, 64, 2, PT_NEE_PAIR)
PT_QUEUE_KERNEL(k_wf_env_pair, 64, 4, PT_ENV_PAIR)
#endif
// phased RIS (LSS_RIS_BSDF_AND_LIGHT, later bounces of LSS_RESTIR_DI): R1 / R2 of wavefront_mis.cuh
PT_QUEUE_KERNEL(k_wf_ris_candidates, 64, 2, ris_candidates_path<PT_GGX>)
PT_QUEUE_KERNEL(k_wf_ris_select, 64, 3, ris_select_path)

// the strategy sample_one_light runs at bounces the phases handle: RIS for LSS 4, and for LSS 5 with RESTIR_DI_LATER_BOUNCES_RIS (bounce >= 1)
#define PT_PHASED_RIS (PT_LSS == 4 || (PT_LSS == 5 && PT_LATER == 3))
__global__ void __launch_bounds__(64, PT_PHASE_BLOCKS_PER_SM) PT_VARIANT_NAME(k_wf_bsdf_resolved)(const __grid_constant__ HIPRTRenderData rd, int2 res, Tile tile, const __grid_constant__ WavefrontState wf, int bounce, bool lights_sampled, bool envmap_sampled)
{
    PT_FOR_EACH_QUEUED(i, wf.counts[kCountsPerBounce * bounce + 1], 7)
        wf_bsdf_resolved_path<PT_ESS, PT_GGX, PT_PHASED_RIS>(rd, res, tile, wf, bounce, i, lights_sampled, envmap_sampled);
}

enum PhaseKernel { PHASE_NEE_LIGHT_EVAL, PHASE_NEE_BSDF_SAMPLE, PHASE_ENV_LIGHT_EVAL, PHASE_ENV_BSDF_SAMPLE, PHASE_BSDF_RESOLVED, PHASE_NEE_PAIR, PHASE_ENV_PAIR, PHASE_RIS_CANDIDATES, PHASE_RIS_SELECT };

cudaError_t PT_VARIANT_NAME(launch_phase)(int which, unsigned blocks, cudaStream_t stream, const HIPRTRenderData* rd, int2 res, Tile tile, const WavefrontState* wf, int bounce,
                                          bool lights_sampled, bool envmap_sampled)
{
    switch (which)
    {
#if PT_LSS == 3
    case PHASE_NEE_LIGHT_EVAL: PT_VARIANT_NAME(k_wf_nee_light_eval)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
    case PHASE_NEE_BSDF_SAMPLE: PT_VARIANT_NAME(k_wf_nee_bsdf_sample)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
#endif
    case PHASE_ENV_LIGHT_EVAL: PT_VARIANT_NAME(k_wf_env_light_eval)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
    case PHASE_ENV_BSDF_SAMPLE: PT_VARIANT_NAME(k_wf_env_bsdf_sample)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
    case PHASE_BSDF_RESOLVED: PT_VARIANT_NAME(k_wf_bsdf_resolved)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce, lights_sampled, envmap_sampled); break;
#if PT_LSS == 3
    case PHASE_NEE_PAIR: PT_VARIANT_NAME(k_wf_nee_pair)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
    case PHASE_ENV_PAIR: PT_VARIANT_NAME(k_wf_env_pair)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
#endif
    case PHASE_RIS_CANDIDATES: PT_VARIANT_NAME(k_wf_ris_candidates)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
    case PHASE_RIS_SELECT: PT_VARIANT_NAME(k_wf_ris_select)<<<blocks, 64, 0, stream>>>(*rd, res, tile, *wf, bounce); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
} // namespace pt
