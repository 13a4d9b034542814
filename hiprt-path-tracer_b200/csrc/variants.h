// Compiled FullPathTracer permutations (LSS, ESS, later-bounces strategy under ReSTIR DI, GGX sampler).
// Values are the reference's KernelOptions.h macro values. One translation unit per permutation
// (b200pt_fpt_variant.cu) so that they compile in parallel. The list is the matrix SURVEY.md section 8(b) asks for:
// {LSS_MIS_LIGHT_BSDF 3, LSS_RIS_BSDF_AND_LIGHT 4, LSS_RESTIR_DI 5} x {ESS_NO_SAMPLING 0, ESS_ALIAS_TABLE 2} x {GGX_VNDF_SAMPLING 0,
// GGX_VNDF_SPHERICAL_CAPS 1} = 12 (reference: src/Compiler/GPUKernelCompilerOptions.cpp:18-70); b200pt_kernel_variant_select refuses the rest.
#pragma once
#define PT_FOR_EACH_VARIANT(M) \
    M(3, 0, 3, 0) M(3, 2, 3, 0) M(3, 2, 3, 1) M(4, 2, 3, 0) M(4, 0, 3, 0) M(5, 0, 3, 0) M(5, 2, 3, 0) \
    M(3, 0, 3, 1) M(4, 0, 3, 1) M(4, 2, 3, 1) M(5, 0, 3, 1) M(5, 2, 3, 1)
