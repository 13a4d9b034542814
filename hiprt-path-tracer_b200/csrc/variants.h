// Compiled FullPathTracer permutations (LSS, ESS, later-bounces strategy under ReSTIR DI, GGX sampler).
// Values are the reference's KernelOptions.h macro values. One translation unit per permutation
// (b200pt_fpt_variant.cu) so that they compile in parallel; the list covers the option sets of configs C1-C5
//  plus the reference's CPU default (RIS + alias table).
#pragma once
#define PT_FOR_EACH_VARIANT(M) \
    M(3, 0, 3, 0) M(3, 2, 3, 0) M(3, 2, 3, 1) M(4, 2, 3, 0) M(4, 0, 3, 0) M(5, 0, 3, 0) M(5, 2, 3, 0)
