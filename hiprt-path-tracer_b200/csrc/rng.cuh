// Per-path random numbers: wang_hash seeding + xorshift32 stream, consumed in the reference's program order.
// reference: src/Device/includes/Hash.h:11-19, src/HostDeviceCommon/Xorshift.h:17-65.
#pragma once

#include "vec.cuh"

namespace pt
{
PT_HD unsigned int wang_hash(unsigned int seed)
{
    seed = (seed ^ 61u) ^ (seed >> 16);
    seed *= 9u;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2du;
    seed = seed ^ (seed >> 15);
    return seed;
}

struct Rng
{
    unsigned int state;

    PT_HD explicit Rng(unsigned int seed) : state(seed) {}

    PT_HD unsigned int next_u32()
    {
        unsigned int x = state;
        x ^= x << 13;
        x ^= x >> 17;
        x ^= x << 5;
        return state = x;
    }

    // u32 / float(0xffffffff) (= 2^32 after rounding), clamped below 1 (Xorshift.h:40-46)
    PT_HD float operator()()
    {
        float a = (float)next_u32() / 4294967296.0f;
        return fminf(a, 1.0f - 1.0e-7f);
    }

    // Xorshift.h:34-38: float product truncated to int, then min with size-1
    PT_HD int random_index(int array_size)
    {
        int r = (int)((float)next_u32() / 4294967296.0f * (float)array_size);
        return r < array_size - 1 ? r : array_size - 1;
    }
};
} // namespace pt
