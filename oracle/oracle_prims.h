/* oracle_prims.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Arithmetic primitives of the CPU oracle, in two interchangeable flavours:
 *
 *  (default)            a from-the-paper restatement of Threefry2x32-20 key
 *                       splitting, Lemire's bounded integers and the handful of
 *                       quaternion/vector formulas the step uses. Travels with
 *                       the repo, needs nothing but a C++17 compiler.
 *  -DMBO_REF_PRIMS      forwards every primitive to the reference's OWN headers
 *                       (external/madrona/include/madrona/rand.hpp, math.hpp),
 *                       included from where they lie under /root/reference. Only
 *                       buildable where the reference is mounted; the binary goes
 *                       to oracle/_ref/. tests/ run both flavours on the same
 *                       seeds and require bit-identical trajectories, which pins
 *                       the restatement to the reference's code.
 *
 * Reference lines restated (M/ = external/madrona/):
 *   RandKey, split_i        M/include/madrona/rand.inl:31-103
 *   initKey                 M/include/madrona/rand.inl:7-10
 *   bits32                  M/include/madrona/rand.inl:105-108
 *   sampleI32 (Lemire)      M/include/madrona/rand.inl:115-159
 *   bitsToFloat01           M/include/madrona/rand.inl:182-196
 *   RNG::advance            M/include/madrona/rand.inl:249-255
 *   Quat::angleAxis         M/include/madrona/math.inl:680-691
 *   Quat operator*          M/include/madrona/math.inl:837-845
 *   Quat::rotateVec         M/include/madrona/math.inl:669-678
 *   Vector3::cross/length/invLength/normalize  M/include/madrona/math.inl:213-220,237-270
 */
#pragma once
#include <cstdint>
#include <cmath>

#ifdef MBO_REF_PRIMS
#include <madrona/rand.hpp>
#include <madrona/math.hpp>
#endif

namespace mbo {

struct Key { uint32_t a, b; };
struct Vec3 { float x, y, z; };
struct Quat { float w, x, y, z; };

#ifdef MBO_REF_PRIMS

static inline Key key_split(Key k, uint32_t idx, uint32_t idx_upper = 0)
{
    madrona::RandKey r = madrona::rand::split_i(madrona::RandKey{k.a, k.b}, idx, idx_upper);
    return Key{r.a, r.b};
}
static inline Key key_init(uint32_t seed)
{
    madrona::RandKey r = madrona::rand::initKey(seed);
    return Key{r.a, r.b};
}
static inline int32_t key_sample_i32(Key k, int32_t a, int32_t b)
{
    return madrona::rand::sampleI32(madrona::RandKey{k.a, k.b}, a, b);
}
static inline float key_sample_uniform(Key k)
{
    return madrona::rand::sampleUniform(madrona::RandKey{k.a, k.b});
}
static inline Quat quat_angle_axis_z(float angle)
{
    madrona::math::Quat q = madrona::math::Quat::angleAxis(angle, madrona::math::Vector3{0.f, 0.f, 1.f});
    return Quat{q.w, q.x, q.y, q.z};
}
static inline Quat quat_mul(Quat a, Quat b)
{
    madrona::math::Quat q = madrona::math::Quat{a.w, a.x, a.y, a.z} * madrona::math::Quat{b.w, b.x, b.y, b.z};
    return Quat{q.w, q.x, q.y, q.z};
}
static inline Vec3 quat_rotate(Quat q, Vec3 v)
{
    madrona::math::Vector3 r = madrona::math::Quat{q.w, q.x, q.y, q.z}.rotateVec(madrona::math::Vector3{v.x, v.y, v.z});
    return Vec3{r.x, r.y, r.z};
}
static inline Vec3 vec_normalize(Vec3 v)
{
    madrona::math::Vector3 r = madrona::math::Vector3{v.x, v.y, v.z}.normalize();
    return Vec3{r.x, r.y, r.z};
}
static inline float vec_length(Vec3 v)
{
    return madrona::math::Vector3{v.x, v.y, v.z}.length();
}
static inline const char *prims_flavour() { return "reference-headers"; }

#else /* restated primitives */

static inline uint32_t rotl32(uint32_t v, unsigned d) { return (v << d) | (v >> (32u - d)); }

/* Threefry-2x32, 20 rounds (Salmon et al., SC'11), keyed by `k`, counter
 * (idx, idx_upper); the JAX-style "split" the reference uses returns the two
 * output words as the new key. */
static inline Key key_split(Key k, uint32_t idx, uint32_t idx_upper = 0)
{
    static const unsigned R[8] = {13, 15, 26, 6, 17, 29, 16, 24};
    const uint32_t ks[3] = {k.a, k.b, 0x1BD11BDAu ^ k.a ^ k.b};
    uint32_t x0 = idx + ks[0];
    uint32_t x1 = idx_upper + ks[1];
    for (unsigned r = 0; r < 20; ++r) {
        x0 += x1;
        x1 = rotl32(x1, R[r & 7]);
        x1 ^= x0;
        if ((r & 3) == 3) {
            unsigned j = r / 4 + 1; /* key-injection number 1..5 */
            x0 += ks[j % 3];
            x1 += ks[(j + 1) % 3] + j;
        }
    }
    return Key{x0, x1};
}

static inline Key key_init(uint32_t seed) { return key_split(Key{seed, 0u}, 0u, 0u); }

static inline uint32_t key_bits32(Key k) { return k.a ^ k.b; }

/* Lemire, "Fast Random Integer Generation in an Interval", algorithm 5; on
 * rejection the reference re-splits the SAMPLE key with index 0 (it does not
 * advance the stream counter). */
static inline int32_t key_sample_i32(Key k, int32_t a, int32_t b)
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

static inline float key_sample_uniform(Key k)
{
    return (float)(key_bits32(k) >> 8) * 0x1p-24f;
}

/* cosf(+-0.05f) / sinf(+-0.05f) as the correctly rounded single-precision
 * values (what glibc and CUDA's cosf/sinf both return for this argument;
 * tests/test_oracle_prims.py checks them against the reference-header build
 * and against float64 math). Quat::angleAxis(angle, z) = (cos(angle/2), 0*s, 0*s, 1*s). */
static inline Quat quat_angle_axis_z(float angle)
{
    const float half = angle / 2.f;
    float c, s;
    if (half == 0.05f) { c = 0x1.ff5c32p-1f; s = 0x1.996deap-5f; }
    else if (half == -0.05f) { c = 0x1.ff5c32p-1f; s = -0x1.996deap-5f; }
    else { c = cosf(half); s = sinf(half); }
    return Quat{c, 0.f * s, 0.f * s, 1.f * s};
}

static inline Quat quat_mul(Quat a, Quat b)
{
    return Quat{
        (a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z),
        (a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y),
        (a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x),
        (a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w),
    };
}

static inline Vec3 vec_cross(Vec3 a, Vec3 o)
{
    return Vec3{a.y * o.z - a.z * o.y, a.z * o.x - a.x * o.z, a.x * o.y - a.y * o.x};
}

/* v + 2 * (cross(p, v) * w + cross(p, cross(p, v))),  p = (x, y, z) */
static inline Vec3 quat_rotate(Quat q, Vec3 v)
{
    Vec3 p{q.x, q.y, q.z};
    Vec3 pv = vec_cross(p, v);
    Vec3 ppv = vec_cross(p, pv);
    Vec3 t{pv.x * q.w + ppv.x, pv.y * q.w + ppv.y, pv.z * q.w + ppv.z};
    /* 2.f * t is implemented by the reference as t * 2.f (operator*(float, Vector3)) */
    return Vec3{v.x + t.x * 2.f, v.y + t.y * 2.f, v.z + t.z * 2.f};
}

static inline float vec_length(Vec3 v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }

/* host branch of Vector3::invLength: 1.f / length() */
static inline Vec3 vec_normalize(Vec3 v)
{
    float inv = 1.f / vec_length(v);
    return Vec3{v.x * inv, v.y * inv, v.z * inv};
}
static inline const char *prims_flavour() { return "restated"; }

#endif

/* RNG (M/include/madrona/rand.hpp:34-56): key + draw counter; every draw is
 * split_i(key, count++) followed by one of the samplers. */
struct Rng {
    Key k{0, 0};
    uint32_t count = 0;
    Key advance() { return key_split(k, count++); }
    int32_t sample_i32(int32_t a, int32_t b) { return key_sample_i32(advance(), a, b); }
    float sample_uniform() { return key_sample_uniform(advance()); }
};

} // namespace mbo
