/* prims_shim.cpp -- TEST INFRASTRUCTURE. Compiles the product's own arithmetic header
 * (madrona-bots_b200/csrc/sim_prims.cuh) for the host with g++ so that CPU tests can compare it
 * bit-for-bit with the oracle's primitives, and checks that the culled sensor (pair_candidates +
 * exact test on candidates) equals the exhaustive sensor. Built by tests/prims_binding.py with
 * -ffp-contract=off (the host equivalent of nvcc -fmad=false). */
#include <stdint.h>
#include <vector>
#include "../madrona-bots_b200/csrc/sim_prims.cuh"

using namespace mbp;

extern "C" {

void shim_split(uint32_t a, uint32_t b, uint32_t idx, uint32_t up, uint32_t out[2])
{
    Key k = key_split(Key{a, b}, idx, up);
    out[0] = k.a; out[1] = k.b;
}
void shim_init_key(uint32_t seed, uint32_t out[2]) { Key k = key_init(seed); out[0] = k.a; out[1] = k.b; }
int32_t shim_sample_i32(uint32_t a, uint32_t b, int32_t lo, int32_t hi) { return key_sample_i32(Key{a, b}, lo, hi); }
float shim_sample_uniform(uint32_t a, uint32_t b) { return key_sample_uniform(Key{a, b}); }

/* rot (w, z) -> turned (w, z) */
void shim_turn(float w, float z, int left, float out[2]) { Rot r = rot_turn(Rot{w, z}, left != 0); out[0] = r.w; out[1] = r.z; }
void shim_forward(float w, float z, float out[2]) { Dir d = rot_forward(Rot{w, z}); out[0] = d.x; out[1] = d.y; }
void shim_normalize(float x, float y, float out[2]) { Dir d = dir_normalize(Dir{x, y}); out[0] = d.x; out[1] = d.y; }
float shim_move(float *x, float *y, float *w, float *z, int f, int b, int l, int r)
{
    Rot rot{*w, *z};
    float len = move_agent(*x, *y, rot, f != 0, b != 0, l != 0, r != 0);
    *w = rot.w; *z = rot.z;
    return len;
}
int shim_chunk_of(float x, float y) { return chunk_of(x, y); }
uint32_t shim_cell_of(float x, float y) { return cell_of(x, y); }
void shim_surroundings(float x, float y, const uint32_t *na, const uint32_t *ts, float out[2])
{
    surroundings(x, y, na, ts, out[0], out[1]);
}
float shim_species_reward(uint32_t c, uint32_t h, uint32_t n0) { return species_reward(c, h, n0); }
float shim_agent_reward(float st, int32_t hp, int ate, int rep, int hit) { return agent_reward(st, hp, ate != 0, rep != 0, hit != 0); }
float shim_view_yaw(float w, float z) { return view_yaw(Rot{w, z}); }

/* exhaustive sensor of one observer: nearest accepted hit per ray, ties to the lowest slot */
static void sense_brute(int n, const float *x, const float *y, int self, float theta, float *t_out, int *slot_out)
{
    for (int r = 0; r < kSensor; ++r) {
        const Ray ray = make_ray(theta, r);
        float best = 0.f; int hit = -1;
        for (int j = 0; j < n; ++j) {
            float t;
            if (ray_sphere(ray, x[self] - x[j], y[self] - y[j], t))
                if (t > kNearSphere && (hit < 0 || t < best)) { best = t; hit = j; }
        }
        t_out[r] = best; slot_out[r] = hit;
    }
}

/* culled sensor: the algorithm of export_kernel, sequentially */
static void sense_culled(int n, const float *x, const float *y, int self, float theta, Dir h, float *t_out,
                         int *slot_out, long long *tests)
{
    Ray rays[kSensor];
    for (int r = 0; r < kSensor; ++r) { rays[r] = make_ray(theta, r); t_out[r] = 0.f; slot_out[r] = -1; }
    for (int j = 0; j < n; ++j) {
        const float ocx = x[self] - x[j], ocy = y[self] - y[j];
        RaySpan sp;
        if (!pair_candidates(ocx, ocy, h, sp)) continue;
        for (int pass = 0; pass < 2; ++pass) {
            const int lo = pass == 0 ? sp.flo : kFwdRays + sp.blo, hi = pass == 0 ? sp.fhi : kFwdRays + sp.bhi;
            for (int r = lo; r <= hi; ++r) {
                float t;
                ++*tests;
                if (ray_sphere(rays[r], ocx, ocy, t))
                    if (t > kNearSphere && (slot_out[r] < 0 || t < t_out[r])) { t_out[r] = t; slot_out[r] = j; }
            }
        }
    }
}

static uint32_t lcg(uint64_t &s) { s = s * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t)(s >> 33); }
static float urand(uint64_t &s) { return (float)(lcg(s) >> 7) * (1.0f / 16777216.0f); }

/* Random worlds (mode 0), worlds crowded into a small patch (mode 1), and targets placed on or
 * next to the tangent of a ray (mode 2). Returns the number of (observer, ray) results that differ. */
long long shim_cull_trials(uint64_t seed, int trials, int n, int mode, long long *exact_tests)
{
    std::vector<float> x(n), y(n);
    long long bad = 0;
    *exact_tests = 0;
    uint64_t s = seed * 2654435761ull + 12345ull;
    for (int tr = 0; tr < trials; ++tr) {
        /* heading as the simulator has it: a (w, z) rotation; yaw and cull heading derive from it */
        const float ang = (urand(s) * 2.f - 1.f) * kPi * 0.5f;
        const float nrm = 1.f + (urand(s) - 0.5f) * 2e-3f;      /* slightly denormalised, as after many turns */
        const Rot rot{cosf(ang) * nrm, sinf(ang) * nrm};
        const float theta = view_yaw(rot);
        const Dir h = cull_heading(rot);
        const float span = mode == 1 ? 6.f : 1.f;
        for (int j = 0; j < n; ++j) {
            x[j] = mode == 1 ? 60.f + urand(s) * span : urand(s) * 127.f;
            y[j] = mode == 1 ? 40.f + urand(s) * span : urand(s) * 95.f;
        }
        if (mode == 2) {
            /* put target 1.. on the tangent line of a random ray of observer 0, +- a few ulps */
            for (int j = 1; j < n; ++j) {
                const int r = (int)(lcg(s) % kSensor);
                const Ray ray = make_ray(theta, r);
                const float dist = 1.2f + urand(s) * 120.f;
                const float off = (urand(s) < 0.5f ? 1.f : -1.f) * (1.f + (urand(s) - 0.5f) * 1e-3f * (urand(s) < 0.5f ? 1.f : 0.01f));
                x[j] = x[0] + ray.dx * dist - ray.dy * off;
                y[j] = y[0] + ray.dy * dist + ray.dx * off;
            }
        }
        float tb[kSensor], tc[kSensor];
        int sb[kSensor], sc[kSensor];
        const int observers = mode == 2 ? 1 : n;
        for (int i = 0; i < observers; ++i) {
            sense_brute(n, x.data(), y.data(), i, theta, tb, sb);
            sense_culled(n, x.data(), y.data(), i, theta, h, tc, sc, exact_tests);
            for (int r = 0; r < kSensor; ++r)
                if (sb[r] != sc[r] || (sb[r] >= 0 && tb[r] != tc[r])) ++bad;
        }
    }
    return bad;
}

/* pair_maybe (stage 1 of the cull) must pass every pair for which pair_run reports candidates. Random pairs
 * over the whole world plus pairs placed next to the wedge edges at every distance. Returns violations;
 * *passed = pairs that survive stage 1. */
long long shim_wedge_trials(uint64_t seed, long long trials, long long *passed)
{
    uint64_t s = seed * 2654435761ull + 99ull;
    long long bad = 0;
    *passed = 0;
    for (long long tr = 0; tr < trials; ++tr) {
        const float ang = (urand(s) * 2.f - 1.f) * kPi * 0.5f;
        const float nrm = 1.f + (urand(s) - 0.5f) * 2e-3f;
        const Rot rot{cosf(ang) * nrm, sinf(ang) * nrm};
        const Dir h = cull_heading(rot);
        float ocx, ocy;
        if (tr & 1) {
            ocx = (urand(s) * 2.f - 1.f) * 127.f; ocy = (urand(s) * 2.f - 1.f) * 95.f;
        } else {
            /* near a wedge edge: direction heading +- pi/4 (+ pi), offset sideways by about one radius */
            const float dist = 1.0f + urand(s) * urand(s) * 160.f;
            const float edge = atan2f(h.y, h.x) + kPi * 0.25f * (float)(1 + 2 * (int)(lcg(s) & 3u));
            const float off = (urand(s) * 2.f - 1.f) * 1.7f;
            ocx = -(cosf(edge) * dist - sinf(edge) * off);
            ocy = -(sinf(edge) * dist + cosf(edge) * off);
        }
        const RayRun rc = pair_run(ocx, ocy, h);
        const bool maybe = pair_maybe(ocx, ocy, h);
        if (maybe) ++*passed;
        if (rc.count > 0 && !maybe) ++bad;
    }
    return bad;
}

} /* extern "C" */
