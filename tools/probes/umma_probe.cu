// umma_probe.cu -- validates a hand-built tcgen05.mma recipe on sm_100a: one 128x128x128 bf16 product,
// A and B K-major in the canonical SWIZZLE_128B shared-memory layout, fp32 accumulator in TMEM, read back
// with tcgen05.ld. Prints the maximum error against a host reference. Round-2 groundwork for a tcgen05
// version of csrc/policy_fused.cu.   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_probe umma_probe.cu
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

constexpr int M = 128, N = 128, K = 128;

__device__ __forceinline__ uint32_t sw128_offset(int row, int k)      // bytes, within a [rows x 128 k] bf16 operand
{
    const int kb = k >> 6, kk = k & 63;
    return (uint32_t)(kb * (128 * 128) + (row >> 3) * 1024 + (row & 7) * 128 + (((kk >> 3) ^ (row & 7)) << 4) + (kk & 7) * 2);
}

__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr)      // K-major, SWIZZLE_128B, 8-row groups 1024 B apart
{
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;            // leading byte offset (16 B units): 1
    d |= (uint64_t)64 << 32;           // stride byte offset: 1024 B
    d |= (uint64_t)1 << 46;            // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;            // SWIZZLE_128B
    return d;
}

__global__ void __launch_bounds__(128) probe(const __nv_bfloat16 *A, const __nv_bfloat16 *B, float *D)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(8) unsigned long long bar;
    unsigned char *sA = smem, *sB = smem + 32768;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < M * K; i += 128) {
        const int r = i / K, k = i % K;
        *reinterpret_cast<__nv_bfloat16 *>(sA + sw128_offset(r, k)) = A[i];
        *reinterpret_cast<__nv_bfloat16 *>(sB + sw128_offset(r, k)) = B[i];
    }
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(&bar);
    if (warp == 0) {
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&tmem_base_s);
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(128u));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar_addr), "r"(1u));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    asm volatile("fence.proxy.async.shared::cta;");        // generic-proxy smem writes -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_base_s;
    if (tid == 0) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t a0 = (uint32_t)__cvta_generic_to_shared(sA), b0 = (uint32_t)__cvta_generic_to_shared(sB);
        for (int k = 0; k < K / 16; ++k) {
            const uint32_t off = (uint32_t)(k >> 2) * 16384u + (uint32_t)(k & 3) * 32u;
            const uint64_t da = make_desc(a0 + off), db = make_desc(b0 + off);
            const uint32_t acc = k > 0 ? 1u : 0u;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                         ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc));
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_addr));
    }
    // everyone waits for the accumulator
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(bar_addr), "r"(0u) : "memory");
        }
    }
    asm volatile("tcgen05.fence::after_thread_sync;");
    for (int c = 0; c < N / 32; ++c) {
        uint32_t v[32];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 32);
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                       "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                       "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const int row = warp * 32 + lane;
        for (int i = 0; i < 32; ++i) D[row * N + c * 32 + i] = __uint_as_float(v[i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u));
}

int main()
{
    std::vector<__nv_bfloat16> hA(M * K), hB(N * K);
    std::vector<float> fA(M * K), fB(N * K), ref(M * N), out(M * N);
    srand(1);
    for (int i = 0; i < M * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hA[i] = __float2bfloat16(v); fA[i] = __bfloat162float(hA[i]); }
    for (int i = 0; i < N * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hB[i] = __float2bfloat16(v); fB[i] = __bfloat162float(hB[i]); }
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) { double s = 0; for (int k = 0; k < K; ++k) s += (double)fA[m * K + k] * fB[n * K + k]; ref[m * N + n] = (float)s; }
    __nv_bfloat16 *dA, *dB; float *dD;
    cudaMalloc(&dA, M * K * 2); cudaMalloc(&dB, N * K * 2); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, hA.data(), M * K * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), N * K * 2, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, M * N * 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 + 1024);
    probe<<<1, 128, 65536 + 1024>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    cudaMemcpy(out.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0, maxref = 0; int bad = 0;
    for (int i = 0; i < M * N; ++i) { double d = fabs(out[i] - ref[i]); if (d > maxerr) maxerr = d; if (fabs(ref[i]) > maxref) maxref = fabs(ref[i]); if (d > 1e-2) ++bad; }
    printf("max |err| = %.6f (max |ref| = %.3f), elements off by > 1e-2: %d of %d\n", maxerr, maxref, bad, M * N);
    printf("D[0][0..3] = %f %f %f %f   ref %f %f %f %f\n", out[0], out[1], out[2], out[3], ref[0], ref[1], ref[2], ref[3]);
    printf("D[77][100] = %f ref %f\n", out[77 * N + 100], ref[77 * N + 100]);
    return bad != 0;
}
