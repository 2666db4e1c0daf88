/* Deterministic single-precision sin/cos shared by host and device.
 *
 * Level generation in the reference draws a yaw angle and builds
 * Quat::angleAxis (M/include/madrona/math.inl:728-739) with libm's
 * cosf/sinf; the resulting AABB gates accept/reject and therefore the RNG
 * stream.  glibc's sinf/cosf and CUDA's are not bit-identical (and glibc's
 * own result depends on the FMA ifunc variant the host CPU selects), so both
 * sides evaluate THIS routine instead: argument reduction and a fixed
 * polynomial in IEEE double using only +,-,* in a fixed order (no FMA
 * contraction: nvcc -fmad=false / gcc -ffp-contract=off), rounded once to
 * float.  The double result is accurate to ~1e-16, so the float is the
 * correctly rounded value except within ~1e-9 ulp of a tie.
 *
 * The CUDA kernels call hs_det_sincosf directly; the oracle build interposes
 * sinf/cosf/sincosf of the reference objects onto it (oracle/ref_stubs.cpp).
 * Polynomial coefficients are the classic fdlibm __kernel_sin/__kernel_cos
 * minimax sets on [-pi/4, pi/4].
 */
#ifndef HIDESEEK_DETMATH_H
#define HIDESEEK_DETMATH_H

#if defined(__CUDACC__)
#define HS_DET_FN __host__ __device__ static inline
#else
#define HS_DET_FN static inline
#endif

#include <math.h>

HS_DET_FN void hs_det_sincos_d(double x, double *s_out, double *c_out)
{
    const double two_over_pi = 6.36619772367581382433e-01;
    const double pio2_hi = 1.57079632673412561417e+00;  /* first 33 bits of pi/2 */
    const double pio2_lo = 6.07710050650619224932e-11;  /* pi/2 - pio2_hi */

    double kd = rint(x * two_over_pi);
    double r = (x - kd * pio2_hi) - kd * pio2_lo;
    long long k = (long long)kd;

    double z = r * r;

    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;

    double ps = S1 + z * (S2 + z * (S3 + z * (S4 + z * (S5 + z * S6))));
    double sr = r + (r * z) * ps;
    double pc = C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6))));
    double cr = (1.0 - 0.5 * z) + (z * z) * pc;

    double s, c;
    switch ((int)(k & 3)) {
    case 0: s = sr; c = cr; break;
    case 1: s = cr; c = -sr; break;
    case 2: s = -sr; c = -cr; break;
    default: s = -cr; c = sr; break;
    }
    *s_out = s;
    *c_out = c;
}

HS_DET_FN void hs_det_sincosf(float x, float *s_out, float *c_out)
{
    double s, c;
    hs_det_sincos_d((double)x, &s, &c);
    *s_out = (float)s;
    *c_out = (float)c;
}

HS_DET_FN float hs_det_sinf(float x)
{
    float s, c;
    hs_det_sincosf(x, &s, &c);
    return s;
}

HS_DET_FN float hs_det_cosf(float x)
{
    float s, c;
    hs_det_sincosf(x, &s, &c);
    return c;
}

#endif
