/* sim_prims.cuh -- arithmetic building blocks of the step kernels.
 *
 * Everything here is __host__ __device__ so that tests/ can compile the very
 * same functions with g++ (tests/prims_shim.cpp) and compare them bit-for-bit
 * with the CPU oracle on the host, without a GPU.
 *
 * The simulation is planar: agents only ever rotate about z and move in the
 * xy-plane (src/sim/sim.cpp:456-475), so a rotation is kept as the (w, z) pair
 * of the reference's quaternion and a direction as (x, y). Each planar formula
 * below is the reference's 3-D formula with the identically-zero terms removed;
 * removing a "+ (+-0)" term never changes a non-zero IEEE result, which is why
 * the (w, z) pair evolves bit-identically to the full quaternion
 * (M/ = external/madrona/):
 *   Quat * Quat            M/include/madrona/math.inl:837-845
 *   Quat::rotateVec        M/include/madrona/math.inl:669-678
 *   Vector3::normalize     M/include/madrona/math.inl:237-270 (host branch: 1/length)
 *   split_i (Threefry)     M/include/madrona/rand.inl:31-103
 *   sampleI32 (Lemire)     M/include/madrona/rand.inl:115-159
 *   bitsToFloat01          M/include/madrona/rand.inl:182-196
 *
 * Build contract: no FMA contraction (nvcc -fmad=false, g++ -ffp-contract=off),
 * IEEE division and square root (nvcc defaults; never -use_fast_math).
 */
#pragma once
#include <stdint.h>
#include <math.h>
#include "../../include/mbots_detmath.h"

#if defined(__CUDACC__)
#define MBP_HD __host__ __device__ __forceinline__
#else
#define MBP_HD static inline
#endif

namespace mbp {

/* ---- constants fixed by the reference (src/entry/mgr.cpp:104-113,132-133) ---- */
constexpr int kNumSpecies = 4;
constexpr int kChunkWidth = 16;
constexpr int kMaxFoodPackages = 5;
constexpr int kChunksX = 8, kChunksY = 6, kNumChunks = 48;
constexpr int kTotalAllowedFood = 30;
constexpr int kSensor = 32;
constexpr int kFwdRays = 24;                 /* 3/4 of the sensor (ecs_system.cpp:181) */
constexpr int kFinderRay = 12;               /* (3*32/4)/2, bvh_raycast.cpp:272 */
constexpr float kNearSphere = 1.1f;
constexpr float kWorldX = 128.f, kWorldY = 96.f;
constexpr float kPi = 3.14159265358979323846264338327950288f;
/* Quat::angleAxis(+-0.1, z) = (cos 0.05, 0, 0, +-sin 0.05), correctly rounded */
constexpr float kRotCos = 0x1.ff5c32p-1f;
constexpr float kRotSin = 0x1.996deap-5f;

/* ------------------------------- RNG ---------------------------------------- */
struct Key { uint32_t a, b; };

MBP_HD uint32_t rotl32(uint32_t v, uint32_t d)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(v, v, d);
#else
    return (v << d) | (v >> (32u - d));
#endif
}

#define MBP_TF_ROUND(rot) do { x0 += x1; x1 = rotl32(x1, rot); x1 ^= x0; } while (0)
#define MBP_TF_QUAD_A() do { MBP_TF_ROUND(13u); MBP_TF_ROUND(15u); MBP_TF_ROUND(26u); MBP_TF_ROUND(6u); } while (0)
#define MBP_TF_QUAD_B() do { MBP_TF_ROUND(17u); MBP_TF_ROUND(29u); MBP_TF_ROUND(16u); MBP_TF_ROUND(24u); } while (0)

/* Threefry-2x32, 20 rounds: block cipher of the counter (idx, idx_upper) under
 * `k`; both output words form the derived key. */
MBP_HD Key key_split(Key k, uint32_t idx, uint32_t idx_upper = 0u)
{
    const uint32_t k0 = k.a, k1 = k.b, k2 = 0x1BD11BDAu ^ k.a ^ k.b;
    uint32_t x0 = idx + k0, x1 = idx_upper + k1;
    MBP_TF_QUAD_A(); x0 += k1; x1 += k2 + 1u;
    MBP_TF_QUAD_B(); x0 += k2; x1 += k0 + 2u;
    MBP_TF_QUAD_A(); x0 += k0; x1 += k1 + 3u;
    MBP_TF_QUAD_B(); x0 += k1; x1 += k2 + 4u;
    MBP_TF_QUAD_A(); x0 += k2; x1 += k0 + 5u;
    return Key{x0, x1};
}

MBP_HD Key key_init(uint32_t seed) { return key_split(Key{seed, 0u}, 0u, 0u); }
MBP_HD uint32_t key_bits32(Key k) { return k.a ^ k.b; }

/* Unbiased integer in [a, b): multiply-shift with rejection. A rejected sample
 * re-keys itself with index 0; the stream counter is not involved. */
MBP_HD int32_t key_sample_i32(Key k, int32_t a, int32_t b)
{
    const uint32_t s = (uint32_t)(b - a);
    uint64_t m = (uint64_t)key_bits32(k) * (uint64_t)s;
    uint32_t lo = (uint32_t)m;
    if (lo < s) {
        const uint32_t t = (0u - s) % s;
        while (lo < t) {
            k = key_split(k, 0u);
            m = (uint64_t)key_bits32(k) * (uint64_t)s;
            lo = (uint32_t)m;
        }
    }
    return (int32_t)(uint32_t)(m >> 32) + a;
}

MBP_HD float key_sample_uniform(Key k) { return (float)(key_bits32(k) >> 8) * 0x1p-24f; }

/* ------------------------------ planar pose --------------------------------- */
struct Rot { float w, z; };
struct Dir { float x, y; };

/* rot * angleAxis(+-0.1, z) */
MBP_HD Rot rot_turn(Rot r, bool left)
{
    const float c = kRotCos, s = left ? kRotSin : -kRotSin;
    Rot o;
    o.w = r.w * c - r.z * s;
    o.z = r.w * s + r.z * c;
    return o;
}

/* rot.rotateVec(x-hat): (1 - 2 z^2, 2 z w), not normalised */
MBP_HD Dir rot_forward(Rot r)
{
    Dir d;
    d.x = 1.f + (-(r.z * r.z)) * 2.f;
    d.y = (r.z * r.w) * 2.f;
    return d;
}

MBP_HD float dir_length(Dir d) { return sqrtf(d.x * d.x + d.y * d.y); }

MBP_HD Dir dir_normalize(Dir d)
{
    const float inv = 1.f / dir_length(d);
    return Dir{d.x * inv, d.y * inv};
}

/* actionSystem movement (src/sim/sim.cpp:456-486): returns |delta pos| */
MBP_HD float move_agent(float &x, float &y, Rot &rot, bool fwd, bool back, bool left, bool right)
{
    if (left) rot = rot_turn(rot, true);
    else if (right) rot = rot_turn(rot, false);
    const float ox = x, oy = y;
    const Dir v = dir_normalize(rot_forward(rot));
    if (fwd) { x += v.x; y += v.y; }
    else if (back) { x -= v.x; y -= v.y; }
    /* std::min(limit - 1, std::max(0, p)) with the std:: tie rules */
    x = (0.f < x) ? x : 0.f;
    x = (x < kWorldX - 1.0f) ? x : kWorldX - 1.0f;
    y = (0.f < y) ? y : 0.f;
    y = (y < kWorldY - 1.0f) ? y : kWorldY - 1.0f;
    return dir_length(Dir{x - ox, y - oy});
}

/* getChunkCoord + getChunkIndex for an in-bounds position (sim.inl:39-58) */
MBP_HD int chunk_of(float x, float y)
{
    const int cx = (int)floorf(x * 0.0625f), cy = (int)floorf(y * 0.0625f);
    return cx + cy * kChunksX;
}

/* cell of a position inside its chunk, as healthSync computes it (sim.cpp:517-523):
 * returns (cell_x << 4) | cell_y */
MBP_HD uint32_t cell_of(float x, float y)
{
    const float chx = x * 0.0625f, chy = y * 0.0625f;
    const uint32_t cx = (uint32_t)(uint8_t)(int)(16.f * (chx - floorf(chx)));
    const uint32_t cy = (uint32_t)(uint8_t)(int)(16.f * (chy - floorf(chy)));
    return (cx << 4) | cy;
}

MBP_HD int chunk_index_checked(float cx, float cy)
{
    const int x = (int)cx, y = (int)cy;
    if (x < 0 || y < 0 || x >= kChunksX || y >= kChunksY) return -1;
    return x + y * kChunksX;
}

/* updateSurroundingObservation (src/sim/sim.cpp:583-654): bilinear interpolation
 * of a per-chunk counter over the four chunk centres around (x, y). `na`/`ts`
 * are the 48-entry numAgents / totalSpeed arrays of the world. */
template <typename A>
MBP_HD void surroundings(float x, float y, const A *na, const A *ts, float &out_n, float &out_s)
{
    const float chx = (x - 8.f) * 0.0625f, chy = (y - 8.f) * 0.0625f;
    const float fx = floorf(chx), fy = floorf(chy), cx = ceilf(chx), cy = ceilf(chy);
    const int i00 = chunk_index_checked(fx, fy), i10 = chunk_index_checked(cx, fy);
    const int i01 = chunk_index_checked(fx, cy), i11 = chunk_index_checked(cx, cy);
    const float xi = chx - fx, yi = chy - fy;
    const float n00 = i00 < 0 ? 0.f : (float)na[i00], n10 = i10 < 0 ? 0.f : (float)na[i10];
    const float n01 = i01 < 0 ? 0.f : (float)na[i01], n11 = i11 < 0 ? 0.f : (float)na[i11];
    const float s00 = i00 < 0 ? 0.f : (float)ts[i00], s10 = i10 < 0 ? 0.f : (float)ts[i10];
    const float s01 = i01 < 0 ? 0.f : (float)ts[i01], s11 = i11 < 0 ? 0.f : (float)ts[i11];
    const float nx0 = xi * n10 + (1.f - xi) * n00;
    const float nx1 = xi * n11 + (1.f - xi) * n01;
    const float sx0 = xi * s10 + (1.f - xi) * s00;
    const float sx1 = xi * s11 + (1.f - xi) * s01;
    out_n = yi * nx1 + (1.f - yi) * nx0;
    out_s = yi * sx1 + (1.f - yi) * sx0;
}

/* speciesInfoSync reward term (src/sim/sim.cpp:809-818) */
MBP_HD float species_reward(uint32_t count, uint32_t health_sum, uint32_t init_agents)
{
    float avg = (float)health_sum / (float)count;
    if (count == 0) avg = 0.f;
    return (float)count / (float)init_agents + avg / 100.0f - 2.f;
}

/* rewardSystem, reward setting 8 (src/sim/sim.cpp:943-956) */
MBP_HD float agent_reward(float species_term, int32_t health, bool ate, bool reproduced, bool hit_enemy)
{
    float r = species_term + (float)health / 100.f - 0.5f;
    if (ate) r += 10.f;
    if (reproduced) r += 10.f;
    if (hit_enemy) r += 15.f;
    return r;
}

/* ------------------------------- sensor ------------------------------------- */
/* camera yaw (M/src/render/ecs_system.cpp:174-178) */
MBP_HD float view_yaw(Rot r)
{
    const Dir f = rot_forward(r);
    return mb_atan2f(f.y, f.x);
}

/* One sensor ray: direction and the per-ray constants of the quadratic
 * (M/src/mw/device/bvh_raycast.cpp:227-250, 102-114, 70-92). */
struct Ray { float dx, dy, a, a4, a2; };

MBP_HD Ray make_ray(float theta, int i)
{
    float rt = theta;
    if (i < kFwdRays) {
        rt = (rt - kPi * 0.25f) + (float)i * (kPi * 0.5f / ((float)kFwdRays - 1.f));
    } else {
        rt += kPi;
        rt = (rt - kPi * 0.25f) + (float)(i - kFwdRays) * (kPi * 0.5f / ((float)(kSensor - kFwdRays) - 1.f));
    }
    float sn, cs;
    mb_sincosf(rt, &sn, &cs);
    const float eps = 0.00001f;
    Ray r;
    r.dx = cs; r.dy = sn;
    if (r.dx == 0.f) r.dx += eps;
    if (r.dy == 0.f) r.dy += eps;
    const float dz = 0.f + eps;                 /* the z component is always 0 -> eps */
    r.a = r.dx * r.dx + r.dy * r.dy + dz * dz;
    r.a4 = 4.f * r.a;
    r.a2 = 2.0f * r.a;
    return r;
}

/* ray vs unit sphere at relative offset oc = origin - centre (oc.z == 0).
 * Returns true and the near root when the discriminant is non-negative. */
MBP_HD bool ray_sphere(const Ray &r, float ocx, float ocy, float &t)
{
    const float bq = 2.0f * (ocx * r.dx + ocy * r.dy);
    const float c = (ocx * ocx + ocy * ocy) - 1.f;
    const float disc = bq * bq - r.a4 * c;
    if (disc < 0.f) return false;
    t = (-bq - sqrtf(disc)) / r.a2;
    return true;
}

/* ---- conservative candidate rays of one (observer, target) pair ------------------
 * The exhaustive sensor tests 32 rays against every other agent of the world. A ray can
 * only produce an ACCEPTED hit on a unit sphere at distance D from the origin when
 *   (1) D > 1.1 - eps   (the near root is never farther than the centre, and t must exceed 1.1), and
 *   (2) the angle between the ray and the direction to the centre is at most asin(1/D),
 * so per pair it is enough to run the exact test (ray_sphere, unchanged) on the few rays whose
 * angle falls inside that window. The window is widened by safety margins that dominate every
 * rounding error involved (heading and ray angles ~1e-6 rad, the float discriminant ~1e-2/D^2),
 * which makes the culled result IDENTICAL to the exhaustive one; tests/test_prims.py checks
 * that on millions of random and near-tangent pairs.
 * Rays 0..23 fan over [-pi/4, pi/4] around the heading, rays 24..31 over the same window
 * around heading + pi. Ranges are inclusive and empty when lo > hi. */
struct RaySpan { int flo, fhi, blo, bhi; };

constexpr float kCullMinDist2 = 1.2f;        /* D < 1.095: t <= D < 1.1, nothing can be accepted */
constexpr float kCullAngleMargin = 5e-4f;    /* rad; covers the 1e-5 rad of cull_atan2 and the approximate division */

/* single-instruction approximations for the cull (MUFU.RSQ / MUFU.RCP with flush-to-zero: without -use_fast_math
 * rsqrtf() and __fdividef() carry range-scaling code around them; the cull's operands are never denormal or huge) */
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float cull_rsqrt(float v) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float cull_rcp(float v) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
#endif

MBP_HD float inv_sqrt_upper(float v)
{
#if defined(__CUDA_ARCH__)
    return cull_rsqrt(v) * 1.0005f;
#else
    return (1.f / sqrtf(v)) * 1.0005f;
#endif
}

/* Unit heading of an observer for the cull (direction of rot * x-hat; the exact yaw the rays
 * use is atan2 of the same vector, so the two agree to ~1e-6 rad). */
MBP_HD Dir cull_heading(Rot r)
{
    const Dir f = rot_forward(r);
#if defined(__CUDA_ARCH__)
    const float inv = cull_rsqrt(f.x * f.x + f.y * f.y);
#else
    const float inv = 1.f / sqrtf(f.x * f.x + f.y * f.y);
#endif
    return Dir{f.x * inv, f.y * inv};
}

/* atan2 to ~1e-5 rad (Abramowitz & Stegun 4.4.49 on the reduced octant); cull use only. */
MBP_HD float cull_atan2(float y, float x)
{
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
#if defined(__CUDA_ARCH__)
    const float q = mn * cull_rcp(mx);
#else
    const float q = mn / mx;
#endif
    const float s = q * q;
    float r = mb_fmaf(0.0208351f, s, -0.0851330f);
    r = mb_fmaf(r, s, 0.1801410f);
    r = mb_fmaf(r, s, -0.3302995f);
    r = mb_fmaf(r, s, 0.9998660f);
    r = r * q;
    if (ay > ax) r = 0x1.921fb6p+0f - r;
    if (x < 0.f) r = 0x1.921fb6p+1f - r;
    return y < 0.f ? -r : r;
}

MBP_HD int cull_ceil(float v)
{
#if defined(__CUDA_ARCH__)
    return __float2int_ru(v);
#else
    return (int)ceilf(v);
#endif
}
MBP_HD int cull_floor(float v)
{
#if defined(__CUDA_ARCH__)
    return __float2int_rd(v);
#else
    return (int)floorf(v);
#endif
}

/* oc = observer - target, h = unit heading of the observer. Cull arithmetic only: explicit FMAs are fine
 * here (nothing below feeds a result, only the choice of which exact tests to run). Straight-line code (the
 * kernel runs it on every lane of a warp): returns the run of candidate rays [lo, lo + count) in sensor
 * numbering (forward fan 0..23, backward fan 24..31); count == 0 for a pair that cannot produce an accepted hit.
 * A target is in front of the observer or behind it, so only one fan is evaluated: a target behind is mirrored
 * through the observer (direction (-along, -left)) and looked up in the backward fan's ray spacing. Only a
 * target closer than 1.5 (angular half-width above pi/4) can reach both fans: all 32 rays are candidates then. */
struct RayRun { int lo, count; };

constexpr float kCullBothFansDist2 = 2.25f;

MBP_HD RayRun pair_run(float ocx, float ocy, Dir h)
{
    const float d2 = mb_fmaf(ocx, ocx, ocy * ocy);
    const bool far_enough = d2 >= kCullMinDist2;                 /* false also for NaN */
    const float x = inv_sqrt_upper(far_enough ? d2 : kCullMinDist2);   /* >= 1/D, <= 0.914 */
    const float half = mb_fmaf(x, mb_fmaf(0.32f * x, x, 1.f), kCullAngleMargin);   /* >= asin(1.0005/D) on (0, 0.914] */
    /* direction to the target (-oc) in the heading frame: along, left */
    const float al = -mb_fmaf(ocx, h.x, ocy * h.y);
    const float lf = mb_fmaf(ocx, h.y, -(ocy * h.x));
    const bool back = al < 0.f;
    const float rho = cull_atan2(back ? -lf : lf, fabsf(al));      /* in [-pi/2, pi/2], about the fan's own axis */
    constexpr float inv_df = ((float)kFwdRays - 1.f) / (kPi * 0.5f);
    constexpr float inv_db = ((float)(kSensor - kFwdRays) - 1.f) / (kPi * 0.5f);
    const float inv_d = back ? inv_db : inv_df;
    const int last = back ? kSensor - kFwdRays - 1 : kFwdRays - 1;
    /* ray k of a fan sits at (k / inv_d - pi/4) from the fan axis: candidates are ceil(rs - hs) .. floor(rs + hs) */
    const float rs = mb_fmaf(rho, inv_d, kPi * 0.25f * inv_d), hs = half * inv_d;
    int lo = cull_ceil(rs - hs), hi = cull_floor(rs + hs);
    lo = lo < 0 ? 0 : lo;
    hi = hi > last ? last : hi;
    RayRun rr;
    rr.lo = back ? kFwdRays + lo : lo;
    rr.count = hi >= lo ? hi - lo + 1 : 0;
    if (d2 < kCullBothFansDist2) { rr.lo = 0; rr.count = kSensor; }
    if (!far_enough) rr.count = 0;
    return rr;
}

/* First, cheap stage of the cull: can the pair have candidates at all? Both fans together cover the double
 * wedge |left| <= |along| around the heading axis. With psi the angle between the target direction and that
 * axis, pair_run() yields candidates only when psi <= pi/4 + half, and
 *   |left| - |along| = D * sqrt(2) * sin(psi - pi/4)  <=  D * sqrt(2) * sin(half)  <  1.54
 * for every D in [1.095, 200] (half >= pi/4 only below D = 1.47, where |left| - |along| <= D anyway). Half of
 * all pairs fail this and never reach pair_run(). */
constexpr float kCullWedgeSlack = 1.6f;

MBP_HD bool pair_maybe(float ocx, float ocy, Dir h)
{
    const float al = mb_fmaf(ocx, h.x, ocy * h.y);
    const float lf = mb_fmaf(ocx, h.y, -(ocy * h.x));
    return fabsf(lf) - fabsf(al) <= kCullWedgeSlack;
}

MBP_HD bool pair_candidates(float ocx, float ocy, Dir h, RaySpan &sp)
{
    sp.flo = sp.blo = 0; sp.fhi = sp.bhi = -1;
    if (!pair_maybe(ocx, ocy, h)) return false;
    const RayRun rr = pair_run(ocx, ocy, h);
    const int end = rr.lo + rr.count;                            /* exclusive */
    if (rr.count > 0 && rr.lo < kFwdRays) { sp.flo = rr.lo; sp.fhi = (end < kFwdRays ? end : kFwdRays) - 1; }
    if (rr.count > 0 && end > kFwdRays) { sp.blo = (rr.lo > kFwdRays ? rr.lo : kFwdRays) - kFwdRays; sp.bhi = end - 1 - kFwdRays; }
    return rr.count > 0;
}

} // namespace mbp
