// Host-compiled view of the product's RNG / math headers (csrc/hs_rng.cuh, hs_math.cuh,
// include/hideseek/detmath.h) for the CPU known-answer tests: the same HS_HD functions the
// kernels call, built with g++ -ffp-contract=off.  Not part of the product libraries.
#include <cstdint>

#include "hs_rng.cuh"
#include <hideseek/detmath.h>

using namespace hs;

extern "C" {

void hk_rand_split(uint32_t a, uint32_t b, uint32_t idx, uint32_t idx_upper, uint32_t *out)
{
    RandKey k = splitKey(RandKey{a, b}, idx, idx_upper);
    out[0] = k.a; out[1] = k.b;
}

void hk_rand_init_key(uint32_t seed, uint32_t *out)
{
    RandKey k = initKey(seed);
    out[0] = k.a; out[1] = k.b;
}

uint32_t hk_rand_bits32(uint32_t a, uint32_t b) { return a ^ b; }

int32_t hk_rand_sample_i32(uint32_t a, uint32_t b, int32_t lo, int32_t hi)
{
    return keySampleI32(RandKey{a, b}, lo, hi);
}

float hk_rand_sample_uniform(uint32_t a, uint32_t b) { return float((a ^ b) >> 8) * 0x1p-24f; }

void hk_rng_stream(uint32_t seed, int n, int32_t lo, int32_t hi, float *uni, int32_t *ints)
{
    Rng rng{initKey(seed), 0};
    for (int i = 0; i < n; i++) {
        uni[i] = rng.sampleUniform();
        ints[i] = rng.sampleI32(lo, hi);
    }
}

// Quat::angleAxis with the deterministic sin/cos, general axis
void hk_quat_angle_axis(float angle, const float *axis, float *out)
{
    float s, c;
    hs_det_sincosf(angle / 2.f, &s, &c);
    out[0] = c; out[1] = axis[0] * s; out[2] = axis[1] * s; out[3] = axis[2] * s;
}

void hk_quat_mul(const float *a, const float *b, float *out)
{
    Quat q = Quat{a[0], a[1], a[2], a[3]} * Quat{b[0], b[1], b[2], b[3]};
    out[0] = q.w; out[1] = q.x; out[2] = q.y; out[3] = q.z;
}

void hk_quat_rotate(const float *q, const float *v, float *out)
{
    Vec3 r = rotate(Quat{q[0], q[1], q[2], q[3]}, Vec3{v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

}
