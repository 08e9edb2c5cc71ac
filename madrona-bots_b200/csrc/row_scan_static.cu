/* row_scan_static.cu -- mbots_row_scan_static: the row scan of the step (tile_sum_kernel + scan_kernel of
 * sim_kernels.cu) as a stand-alone entry point that ALLOCATES NOTHING.
 *
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The scan's only scratch
 * memory -- one TileSum per tile of 1,024 worlds -- lives in a file-scope __device__ array of fixed size instead of
 * the handle's cudaMalloc'd tileSum buffers. Every GPU that loads the library holds the array (8 KiB) until exit.
 *
 * The array is SHARED by all calls on a device and keeps the last call's contents. Nothing in it has to start at
 * zero: the first kernel overwrites every tile entry the second one reads. Two calls that overlap on the GPU (different
 * streams, no dependency between them) race on it; serialise them.
 *
 * The existing code paths (mbots_step* and the handle's own scan) do not use anything in this file.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>

#include "../../include/mbots.h"
#include "sim_dev.h"

namespace mbk { int set_error(int code, const char *what, cudaError_t ce); }   /* mbots_api.cu: feeds mbots_last_error */

using namespace mbk;

namespace {

constexpr unsigned FULLMASK = 0xffffffffu;
constexpr uint32_t kStaticMaxTiles = MBOTS_ROW_SCAN_STATIC_MAX_WORLDS / kScanTile;
static_assert(MBOTS_ROW_SCAN_STATIC_MAX_WORLDS % kScanTile == 0, "whole tiles");
static_assert(sizeof(TileSum) == 32, "scratch size below assumes 32-byte tile records");

/* THE scratch: 256 tiles x 32 B = 8 KiB per GPU. In host code this name is a host-side stand-in, not a device
 * address: the host passes what cudaGetSymbolAddress returns for the current device (scratch_for_device below). */
static __device__ __align__(256) TileSum g_rowScanScratch[kStaticMaxTiles];

/* per-tile sums of the (world, species) counts and the largest world of the tile: one CTA per tile */
__global__ void __launch_bounds__(kScanTile) tile_sum_static_kernel(const int4 *__restrict__ live, int W, TileSum *scratch)
{
    __shared__ int red[32][5];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int w = blockIdx.x * kScanTile + tid;
    int4 c = make_int4(0, 0, 0, 0);
    if (w < W) c = live[w];
    const int v[5] = {c.x, c.y, c.z, c.w, c.x + c.y + c.z + c.w};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int r = k == 4 ? __reduce_max_sync(FULLMASK, v[k]) : __reduce_add_sync(FULLMASK, v[k]);
        if (lane == 0) red[warp][k] = r;
    }
    __syncthreads();
    if (tid < 5) {
        int a = 0;
        for (int i = 0; i < kScanTile / 32; ++i) a = tid == 4 ? max(a, red[i][tid]) : a + red[i][tid];
        TileSum *ts = scratch + blockIdx.x;
        if (tid == 4) ts->maxCount = a; else ts->c[tid] = a;
    }
}

__device__ __forceinline__ int incl_scan32(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(FULLMASK, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

/* Rows are ordered (species, world): row_base[w][k] = rows of species < k (all worlds) + rows of species k in worlds
 * < w. Every CTA reduces the tile records to the species totals and the prefix of its own tile, then scans its worlds. */
__global__ void __launch_bounds__(kScanTile) row_scan_static_kernel(const int4 *__restrict__ live, int W,
                                                                    const TileSum *scratch, int4 *row_base,
                                                                    int32_t *world_offset, uint32_t *totals)
{
    __shared__ int red[32][9];
    __shared__ int fin[9];
    __shared__ int wsum[32][5];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, nt = gridDim.x;

    int acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};          /* totals[4], prefix of this tile[4], largest world */
    for (int t = tid; t < nt; t += kScanTile) {
        const TileSum x = scratch[t];
#pragma unroll
        for (int k = 0; k < 4; ++k) { acc[k] += x.c[k]; if (t < tile) acc[4 + k] += x.c[k]; }
        acc[8] = max(acc[8], x.maxCount);
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const int r = k == 8 ? __reduce_max_sync(FULLMASK, acc[k]) : __reduce_add_sync(FULLMASK, acc[k]);
        if (lane == 0) red[warp][k] = r;
    }
    const int w = tile * kScanTile + tid;
    int4 c = make_int4(0, 0, 0, 0);
    if (w < W) c = live[w];
    const int v[5] = {c.x, c.y, c.z, c.w, c.x + c.y + c.z + c.w};
    int inc[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        inc[k] = incl_scan32(v[k], lane);
        if (lane == 31) wsum[warp][k] = inc[k];
    }
    __syncthreads();
    if (tid < 9) {
        int a = 0;
        for (int i = 0; i < kScanTile / 32; ++i) a = tid == 8 ? max(a, red[i][tid]) : a + red[i][tid];
        fin[tid] = a;
    }
    __syncthreads();
    int base[5], species_start = 0, worlds_before = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        base[k] = species_start + fin[4 + k];
        species_start += fin[k];
        worlds_before += fin[4 + k];
    }
    base[4] = worlds_before;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        int b = base[k];
        for (int i = 0; i < warp; ++i) b += wsum[i][k];
        base[k] = b + inc[k] - v[k];
    }
    if (w < W) {
        row_base[w] = make_int4(base[0], base[1], base[2], base[3]);
        if (world_offset) world_offset[w] = base[4];
    }
    if (tile == 0 && tid == 0 && totals) {
        for (int k = 0; k < 4; ++k) totals[k] = (uint32_t)fin[k];
        totals[4] = (uint32_t)species_start;
        totals[5] = (uint32_t)fin[8];
    }
}

/* Device address of g_rowScanScratch on device `dev` (which must be current): looked up once per device. */
constexpr int kMaxDevices = 64;
std::mutex g_setupLock;
TileSum *g_scratchAddr[kMaxDevices];

cudaError_t scratch_for_device(int dev, TileSum **out)
{
    std::lock_guard<std::mutex> hold(g_setupLock);
    if (g_scratchAddr[dev] == nullptr) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_rowScanScratch);
        if (e != cudaSuccess) return e;
        if (p == nullptr) return cudaErrorInvalidSymbol;
        g_scratchAddr[dev] = (TileSum *)p;
    }
    *out = g_scratchAddr[dev];
    return cudaSuccess;
}

/* the device that owns `p`, or -1 when it is not device memory */
int device_of(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return -1; }
    if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged) return -1;
    return a.device;
}

struct CurrentDevice {
    int prev = -1;
    bool ok = false;
    explicit CurrentDevice(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) return;
        ok = prev == dev || cudaSetDevice(dev) == cudaSuccess;
        if (prev == dev) prev = -1;
    }
    ~CurrentDevice() { if (prev >= 0) cudaSetDevice(prev); }
};

} // namespace

extern "C" int mbots_row_scan_static(const int32_t *species_rows, uint32_t num_worlds, int32_t *row_base,
                                     int32_t *world_offset, uint32_t *totals, void *stream)
{
    if (!species_rows || !row_base) return set_error(MBOTS_EINVAL, "null argument", cudaSuccess);
    if (num_worlds == 0) return set_error(MBOTS_EINVAL, "num_worlds must be > 0", cudaSuccess);
    if (num_worlds > MBOTS_ROW_SCAN_STATIC_MAX_WORLDS)
        return set_error(MBOTS_EINVAL,
                         "mbots_row_scan_static: num_worlds exceeds MBOTS_ROW_SCAN_STATIC_MAX_WORLDS (262,144), the size of "
                         "the library's static scratch; use a simulator handle (mbots_step) for larger populations",
                         cudaSuccess);
    if (((uintptr_t)species_rows | (uintptr_t)row_base) & 15u)
        return set_error(MBOTS_EINVAL, "species_rows and row_base must be 16-byte aligned", cudaSuccess);
    const int dev = device_of(species_rows);
    if (dev < 0 || dev >= kMaxDevices || device_of(row_base) != dev || (world_offset && device_of(world_offset) != dev) ||
        (totals && device_of(totals) != dev))
        return set_error(MBOTS_EINVAL, "every buffer must be device memory of one GPU", cudaSuccess);

    CurrentDevice cur(dev);
    if (!cur.ok) return set_error(MBOTS_ECUDA, "cudaSetDevice", cudaSuccess);
    TileSum *scratch = nullptr;
    cudaError_t e = scratch_for_device(dev, &scratch);
    if (e != cudaSuccess) return set_error(MBOTS_ECUDA, "cudaGetSymbolAddress(g_rowScanScratch)", e);

    const unsigned tiles = (num_worlds + kScanTile - 1) / kScanTile;
    cudaStream_t s = (cudaStream_t)stream;
    tile_sum_static_kernel<<<tiles, kScanTile, 0, s>>>((const int4 *)species_rows, (int)num_worlds, scratch);
    row_scan_static_kernel<<<tiles, kScanTile, 0, s>>>((const int4 *)species_rows, (int)num_worlds, scratch,
                                                       (int4 *)row_base, world_offset, totals);
    e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(MBOTS_ECUDA, "mbots_row_scan_static launch", e);
    return MBOTS_OK;
}
