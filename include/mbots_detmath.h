/* mbots_detmath.h -- deterministic single-precision transcendentals.
 *
 * The reference evaluates the sensor ray directions with cosf/sinf
 * (external/madrona/src/mw/device/bvh_raycast.cpp:250) and the camera yaw
 * with atan2f (external/madrona/src/render/ecs_system.cpp:178). Those calls
 * resolve to a different implementation on every platform (CUDA libdevice,
 * -use_fast_math intrinsics, glibc), and their last-ulp differences feed the
 * integer sensor outputs (uint8 depth, hit / no-hit, finder target).
 *
 * To make "bit-exact on integer state" a meaningful, testable contract the
 * simulator DEFINES these three functions here, using only operations that
 * are correctly rounded on both the host and sm_100a (+, -, *, /, fma,
 * rint). The CUDA kernels and the CPU oracle both include this header, so
 * they agree bit-for-bit by construction; tests/test_detmath.py pins the
 * functions themselves against libm (<= 2 ulp over the ranges the simulator
 * uses) so a bug here cannot hide as a common-mode error.
 *
 * Build requirements for bit equality:
 *   host:   -ffp-contract=off   (no implicit FMA contraction)
 *   device: -fmad=false         (same)
 * Explicit mb_fmaf() calls are real fused multiply-adds on both sides.
 */
#ifndef MBOTS_DETMATH_H
#define MBOTS_DETMATH_H

#if defined(__CUDACC__)
#define MB_HD __host__ __device__ __forceinline__
#else
#include <math.h>
#define MB_HD static inline
#endif

MB_HD float mb_fmaf(float a, float b, float c)
{
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}

MB_HD float mb_rintf(float a)
{
#if defined(__CUDA_ARCH__)
    return rintf(a); /* cvt.rni.f32.f32 */
#else
    return rintf(a); /* default rounding mode: nearest-even */
#endif
}

/* sin and cos of x for |x| < ~100 (the simulator uses |x| < 8).
 * Cody-Waite reduction by pi/2 in three fma steps, then degree-7 / degree-8
 * minimax polynomials on [-pi/4, pi/4]. */
MB_HD void mb_sincosf(float x, float *s_out, float *c_out)
{
    const float two_over_pi = 0x1.45f306p-1f;   /* fl32(2/pi) */
    const float pio2_hi = 0x1.921fb6p+0f;       /* fl32(pi/2) */
    const float pio2_mid = -0x1.777a5cp-25f;    /* fl32(pi/2 - hi) */
    const float pio2_lo = -0x1.ee59dap-50f;     /* fl32(pi/2 - hi - mid) */

    float k = mb_rintf(x * two_over_pi);
    float r = mb_fmaf(-k, pio2_hi, x);
    r = mb_fmaf(-k, pio2_mid, r);
    r = mb_fmaf(-k, pio2_lo, r);
    float z = r * r;

    /* sin(r) = r + r*z*(S1 + z*(S2 + z*S3)) */
    float sp = mb_fmaf(-1.95152959e-4f, z, 8.33216087e-3f);
    sp = mb_fmaf(sp, z, -1.66666546e-1f);
    float s = mb_fmaf(r * z, sp, r);

    /* cos(r) = 1 - z/2 + z*z*(C1 + z*(C2 + z*C3)) */
    float cp = mb_fmaf(2.44331571e-5f, z, -1.38873163e-3f);
    cp = mb_fmaf(cp, z, 4.16666457e-2f);
    float c = mb_fmaf(z * z, cp, mb_fmaf(-0.5f, z, 1.0f));

    int q = (int)k & 3;
    float sv = (q & 1) ? c : s;
    float cv = (q & 1) ? s : c;
    if (q == 1 || q == 2) cv = -cv;
    if (q >= 2) sv = -sv;
    *s_out = sv;
    *c_out = cv;
}

/* atan2f(y, x), result in (-pi, pi]. Not both zero. */
MB_HD float mb_atan2f(float y, float x)
{
    const float pi = 0x1.921fb6p+1f;   /* fl32(pi) */
    const float pio2 = 0x1.921fb6p+0f; /* fl32(pi/2) */
    float ax = x < 0.0f ? -x : x;
    float ay = y < 0.0f ? -y : y;
    float mx = ax > ay ? ax : ay;
    float mn = ax > ay ? ay : ax;
    float a = (mx == 0.0f) ? 0.0f : mn / mx;
    float s = a * a;
    float r = 0.0027856871f;
    r = mb_fmaf(r, s, -0.0158660002f);
    r = mb_fmaf(r, s, 0.0424722210f);
    r = mb_fmaf(r, s, -0.0749753043f);
    r = mb_fmaf(r, s, 0.1064487990f);
    r = mb_fmaf(r, s, -0.1420703080f);
    r = mb_fmaf(r, s, 0.1999345420f);
    r = mb_fmaf(r, s, -0.3333314660f);
    r = mb_fmaf(r * s, a, a);
    if (ay > ax) r = pio2 - r;
    if (x < 0.0f) r = pi - r;
    if (y < 0.0f) r = -r;
    return r;
}

#endif /* MBOTS_DETMATH_H */
