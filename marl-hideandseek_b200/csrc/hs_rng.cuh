// Counter-based RNG: Threefry2x32-20 key splitting, Lemire unbiased integers,
// 24-bit uniform floats.  Integer-exact restatement of the published
// Threefry2x32 algorithm as used by the reference
// (M/include/madrona/rand.inl:31-103 split_i, 115-159 sampleI32, 196-210
// bitsToFloat01, 269-275 RNG::advance).
#pragma once

#include <cstdint>
#include "hs_math.cuh"

namespace hs {

struct RandKey { uint32_t a, b; };

HS_HD uint32_t rotl32(uint32_t v, uint32_t d) { return (v << d) | (v >> (32 - d)); }

HS_HD RandKey splitKey(RandKey src, uint32_t idx, uint32_t idx_upper = 0)
{
    const uint32_t ks0 = src.a, ks1 = src.b;
    const uint32_t ks2 = 0x1BD11BDAu ^ src.a ^ src.b;
    uint32_t x0 = idx + ks0;
    uint32_t x1 = idx_upper + ks1;

#define HS_TF_ROUND(r) { x0 += x1; x1 = rotl32(x1, r); x1 ^= x0; }
#define HS_TF_R4A HS_TF_ROUND(13) HS_TF_ROUND(15) HS_TF_ROUND(26) HS_TF_ROUND(6)
#define HS_TF_R4B HS_TF_ROUND(17) HS_TF_ROUND(29) HS_TF_ROUND(16) HS_TF_ROUND(24)
    HS_TF_R4A
    x0 += ks1; x1 += ks2 + 1u;
    HS_TF_R4B
    x0 += ks2; x1 += ks0 + 2u;
    HS_TF_R4A
    x0 += ks0; x1 += ks1 + 3u;
    HS_TF_R4B
    x0 += ks1; x1 += ks2 + 4u;
    HS_TF_R4A
    x0 += ks2; x1 += ks0 + 5u;
#undef HS_TF_R4A
#undef HS_TF_R4B
#undef HS_TF_ROUND
    return {x0, x1};
}

HS_HD RandKey initKey(uint32_t seed) { return splitKey(RandKey{seed, 0}, 0); }

HS_HD int32_t keySampleI32(RandKey k, int32_t a, int32_t b)
{
    uint32_t s = (uint32_t)(b - a);
    uint32_t x = k.a ^ k.b;
    uint64_t m = (uint64_t)x * (uint64_t)s;
    uint32_t l = (uint32_t)m, h = (uint32_t)(m >> 32);
    if (l < s) {
        uint32_t t = (0u - s) % s;
        while (l < t) {
            k = splitKey(k, 0);
            x = k.a ^ k.b;
            m = (uint64_t)x * (uint64_t)s;
            l = (uint32_t)m;
            h = (uint32_t)(m >> 32);
        }
    }
    return (int32_t)h + a;
}

struct Rng {
    RandKey k;
    uint32_t count;

    HS_HD RandKey advance() { RandKey s = splitKey(k, count); count += 1; return s; }
    HS_HD int32_t sampleI32(int32_t a, int32_t b) { return keySampleI32(advance(), a, b); }
    HS_HD float sampleUniform()
    {
        RandKey s = advance();
        return float((s.a ^ s.b) >> 8) * 0x1p-24f;
    }
};

}  // namespace hs
