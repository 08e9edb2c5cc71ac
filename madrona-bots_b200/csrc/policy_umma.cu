/* policy_umma.cu -- the acting path of learn/training_loop.py:52-83,136 on the 5th-generation tensor cores.
 *
 * Same contract as policy_fused.cu (one kernel per species: construct_obs -> species network -> softmax ->
 * Threefry sample -> one-hot action row), different engine: every layer is a 128 x N x K `tcgen05.mma` issued by
 * ONE thread, operands in shared memory in the canonical K-major SWIZZLE_128B layout, accumulator in tensor
 * memory. A persistent CTA per SM keeps TWO tiles of 128 agents in flight: while the eight epilogue warps of one
 * tile pull its accumulator out of TMEM (tcgen05.ld: one thread = one agent row), apply bias / activation / gate
 * arithmetic and write the next layer's A operand back to shared memory (bf16, swizzled), the tensor core is
 * already multiplying the other tile. Weight blocks (bf16, pre-swizzled by the host, 32 KB each) arrive through a
 * three-deep ring of 1-D bulk copies (cp.async.bulk -> mbarrier complete_tx) issued by the MMA thread.
 *
 *   warps 0-7   epilogue of tile 0        warps 8-15  epilogue of tile 1        warp 16  lane 0: copies + MMA issue
 *               (warp & 3 = TMEM lane quarter = 32 agent rows, (warp >> 2) & 1 = which 64 of the 128 columns)
 *   mbarriers   w_full[3] / w_empty[3]    weight ring
 *               acc_full[2]               tcgen05.commit: accumulator of tile t is complete
 *               a_ready[2]                epilogue of tile t drained TMEM and (re)wrote its A operand
 * The descriptor recipe was validated in isolation first (tools/probes/umma_probe.cu).
 */
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "sim_dev.h"
#include "policy_dev.cuh"
#include "../../include/mbots_policy.h"

namespace mbk {

constexpr int kUTile = 128;                         /* agents per tile = UMMA M */
constexpr int kUBlockBytes = 128 * 128 * 2;          /* one operand: [128 rows][128 k] bf16 = two 16 KB K-blocks */
constexpr int kUEpiWarps = 16;                      /* per tile: 4 TMEM lane quarters x 2 column halves */
constexpr int kUThreads = (kUEpiWarps + 1) * 32;    /* + 1 issue warp */
constexpr int kURing = 3;
constexpr int kUTmemCols = 512;                     /* per tile: 128 accumulator columns + 128 scratch columns (x2 tiles, padded to 256 each) */

/* shared memory: [weights x3][A operand x2] + barriers (1 KB aligned base) */
constexpr int kUBiasFloats = MBOTS_POLICY_MAX_OPS * 512;      /* every bias vector of the program, staged once per CTA */
constexpr int kUSmemBytes = (kURing + 2) * kUBlockBytes + 1024 /* alignment slack */ + 256 /* barriers */ + kUBiasFloats * 4;
size_t policy_umma_smem_bytes() { return kUSmemBytes; }

/* byte offset of (row, k) in a K-major SWIZZLE_128B operand: K-blocks of 64, 8-row atoms of 1024 B, 16-byte
 * chunks XOR-ed with the row inside the atom */
__host__ __device__ __forceinline__ uint32_t sw128_chunk_offset(uint32_t row, uint32_t chunk /* k / 8 */)
{
    return (chunk >> 3) * 16384u + (row >> 3) * 1024u + (row & 7u) * 128u + (((chunk & 7u) ^ (row & 7u)) << 4);
}

__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)1 << 16) /* LBO 16 B */ | ((uint64_t)64 << 32) /* SBO 1024 B */ |
           ((uint64_t)1 << 46) /* version */ | ((uint64_t)2 << 61) /* SWIZZLE_128B */;
}

__device__ __forceinline__ uint32_t umma_idesc(uint32_t n)      /* bf16 x bf16 -> f32, M = 128, both K-major */
{
    return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((128u >> 4) << 24);
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

/* 32 consecutive accumulator columns of this thread's row (TMEM lane) */
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void sts128(uint32_t addr, uint4 v)
{
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t upack_bf16(float lo, float hi)
{
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
}
__device__ __forceinline__ float2 uunpack_bf16(uint32_t u)
{
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&u));
}

__device__ __forceinline__ float utanh(float x)
{
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float usigmoid(float x) { return fmaf(0.5f, utanh(0.5f * x), 0.5f); }

template <int ACT>
__device__ __forceinline__ float uact(float x)
{
    if (ACT == MBOTS_ACT_TANH) return utanh(x);
    if (ACT == MBOTS_ACT_ELU) return x > 0.f ? x : __expf(x) - 1.f;
    if (ACT == MBOTS_ACT_LOGSIGMOID) return fminf(x, 0.f) - __logf(1.f + __expf(-fabsf(x)));
    if (ACT == MBOTS_ACT_LEAKYRELU) return x > 0.f ? x : 0.01f * x;
    if (ACT == MBOTS_ACT_RELU) return fmaxf(x, 0.f);
    if (ACT == MBOTS_ACT_SIGMOID) return usigmoid(x);
    return x;
}

/* 16 consecutive columns of this thread's TMEM lane; asynchronous until the matching wait. The waits take the
 * destination registers as in/out operands so that no use of them can be scheduled ahead of the wait. */
#define U16(v, q) q(v[0]), q(v[1]), q(v[2]), q(v[3]), q(v[4]), q(v[5]), q(v[6]), q(v[7]), q(v[8]), q(v[9]), q(v[10]), q(v[11]), q(v[12]), q(v[13]), q(v[14]), q(v[15])
#define UQ_OUT(x) "=r"(x)
#define UQ_IO(x) "+r"(x)
#define UQ_IN(x) "r"(x)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : U16(v, UQ_OUT) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld(uint32_t (&v)[16])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;" : U16(v, UQ_IO) :: "memory");
}
__device__ __forceinline__ void tmem_wait_ld(uint32_t (&v)[16], uint32_t (&s)[16])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;" : U16(v, UQ_IO), U16(s, UQ_IO) :: "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), U16(v, UQ_IN) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ bool elect_one()
{
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}

__device__ __forceinline__ void load_bias16(uint32_t bias /* shared-space address */, float (&b)[16])
{
#pragma unroll
    for (int j = 0; j < 4; ++j)
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(b[4 * j]), "=f"(b[4 * j + 1]), "=f"(b[4 * j + 2]), "=f"(b[4 * j + 3]) : "r"(bias + 16u * j));
}

/* One weight block's epilogue for one thread: its row, 64 of the 128 columns, in four chunks of 16. The TMEM loads
 * of chunk c + 1 are in flight while chunk c is computed; TMEM stores are waited for once, at the end. */
enum { EM_LINEAR = 0, EM_LSTM_I, EM_LSTM_G, EM_LSTM_O, EM_GRU_R, EM_GRU_Z, EM_GRU_N };

template <int MODE, int ACT>
__device__ __forceinline__ void epi_block(uint32_t acc, uint32_t scratch, uint32_t bias, uint32_t bias_hn,
                                          uint32_t a_tile, uint32_t row, int col_base)
{
    constexpr bool kLoadScr = MODE == EM_LSTM_G || MODE == EM_LSTM_O || MODE == EM_GRU_Z || MODE == EM_GRU_N;
    constexpr bool kStoreScr = MODE == EM_LSTM_I || MODE == EM_LSTM_G || MODE == EM_GRU_R || MODE == EM_GRU_Z;
    uint32_t v[2][16], s[2][16];
    float bb[2][16];
    tmem_ld16(acc + (uint32_t)col_base, v[0]);
    if (kLoadScr) tmem_ld16(scratch + (uint32_t)col_base, s[0]);
    load_bias16(bias + 4u * col_base, bb[0]);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const int cur = c & 1, col = col_base + c * 16;
        const float (&b)[16] = bb[cur];
        if (kLoadScr) tmem_wait_ld(v[cur], s[cur]); else tmem_wait_ld(v[cur]);
        if (c + 1 < 4) {
            tmem_ld16(acc + (uint32_t)(col + 16), v[cur ^ 1]);
            if (kLoadScr) tmem_ld16(scratch + (uint32_t)(col + 16), s[cur ^ 1]);
            load_bias16(bias + 4u * (col + 16), bb[cur ^ 1]);
        }
        float x[16];
        uint32_t out[16];
        if (MODE == EM_GRU_R) {                                   /* park r * b_hn (low bf16 half) */
            float hn[16];
            load_bias16(bias_hn + 4u * col, hn);
#pragma unroll
            for (int i = 0; i < 16; ++i) out[i] = upack_bf16(usigmoid(__uint_as_float(v[cur][i]) + b[i]) * hn[i], 0.f);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float pre = __uint_as_float(v[cur][i]) + b[i];
            if (MODE == EM_LINEAR) x[i] = uact<ACT>(pre);
            if (MODE == EM_LSTM_I) out[i] = __float_as_uint(usigmoid(pre));                                   /* sigmoid(i) */
            if (MODE == EM_LSTM_G) out[i] = __float_as_uint(__uint_as_float(s[cur][i]) * utanh(pre));         /* c = sigmoid(i) tanh(g) */
            if (MODE == EM_LSTM_O) x[i] = usigmoid(pre) * utanh(__uint_as_float(s[cur][i]));                  /* h = sigmoid(o) tanh(c) */
            if (MODE == EM_GRU_Z) out[i] = upack_bf16(uunpack_bf16(s[cur][i]).x, 1.f - usigmoid(pre));        /* add 1 - z (high half) */
            if (MODE == EM_GRU_N) {                                                                          /* h = (1 - z) tanh(n + b_in + r b_hn) */
                const float2 rz = uunpack_bf16(s[cur][i]);
                x[i] = rz.y * utanh(pre + rz.x);
            }
        }
        if (kStoreScr) {
            tmem_st16(scratch + (uint32_t)col, out);
        } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const uint4 pk = make_uint4(upack_bf16(x[8 * j], x[8 * j + 1]), upack_bf16(x[8 * j + 2], x[8 * j + 3]),
                                            upack_bf16(x[8 * j + 4], x[8 * j + 5]), upack_bf16(x[8 * j + 6], x[8 * j + 7]));
                sts128(a_tile + sw128_chunk_offset(row, (uint32_t)((col >> 3) + j)), pk);
            }
        }
    }
    if (kStoreScr) tmem_wait_st();
}

template <int KS>
__device__ __forceinline__ void issue_mmas(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc)
{
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {                 /* k-step of 16: +32 B inside a 64-wide K-block, +16 KB to the next one */
        const uint64_t off = (uint64_t)(((ks >> 2) * 16384 + (ks & 3) * 32) >> 4);
        umma_bf16(tmem_d, adesc + off, bdesc + off, idesc, ks > 0 ? 1u : 0u);
    }
}

#ifdef MBK_UMMA_TRACE
/* debug build only: a clock64 timeline of CTA 0 (tools/umma_trace.py); five actors (issuer, 4 epilogue lanes),
 * each with its own region and a register counter: plain stores, no round trips */
__device__ unsigned long long g_umma_trace[5 * 2 * 2048];
#define UTRACE_DECL(actor_) uint32_t tr_n = 0; const uint32_t tr_actor = (actor_);
#define UTRACE(code, t, b)                                                                                      \
    if (blockIdx.x == 0 && tr_n < 2048) {                                                                       \
        g_umma_trace[(tr_actor * 2048 + tr_n) * 2] = (code) | ((t) << 8) | ((uint32_t)(b) << 16) | 0x80000000u;  \
        g_umma_trace[(tr_actor * 2048 + tr_n) * 2 + 1] = clock64();                                             \
        ++tr_n;                                                                                                 \
    }
#else
#define UTRACE_DECL(actor_)
#define UTRACE(code, t, b)
#endif

constexpr int kUMaxNets = 4;                       /* one network per species */

struct UmmaNet {                                   /* one species network: weight blocks (32 KB each) in streaming order */
    const unsigned char *blob;
    int32_t n_blocks;
    int32_t n_ops;
    uint32_t ofs[3 * MBOTS_POLICY_MAX_OPS];
    uint8_t op_of[3 * MBOTS_POLICY_MAX_OPS];
    uint8_t gate_of[3 * MBOTS_POLICY_MAX_OPS];
    uint8_t opcode[MBOTS_POLICY_MAX_OPS];
    uint8_t act[MBOTS_POLICY_MAX_OPS];
    uint16_t bias_at[MBOTS_POLICY_MAX_OPS];        /* float offset of each op's biases in the shared-memory copy */
    uint16_t bias_n[MBOTS_POLICY_MAX_OPS];
    uint32_t b_ofs[MBOTS_POLICY_MAX_OPS];
    uint32_t row_begin, rows;                      /* the rows it acts on, unless they are resolved on the device */
};

/* What one launch does: n networks, each over its own block of rows. species0 >= 0: network j acts on the block of species
 * species0 + j of the current row set, read from `totals` on the device. */
struct UmmaWork {
    int32_t n;
    int32_t species0;
    const Totals *totals;
    UmmaNet net[kUMaxNets];
};

/* the launch's schedule, worked out once per CTA (shared memory): tile pairs are numbered through all networks */
struct UmmaSched {
    uint32_t row_begin[kUMaxNets], rows[kUMaxNets], pair_end[kUMaxNets], out_base[kUMaxNets];
    uint32_t pairs;
};

struct UObsPtrs { const uint8_t *depth; const float *health; const float *position; const int8_t *semantic; const float *surrounding;
                  const uint32_t *owner; uint32_t worldOff; };

__global__ void __launch_bounds__(kUThreads, 1)     /* 17 warps are allocated as 20: 96 registers per thread is the ceiling */
policy_umma_kernel(UObsPtrs o, int32_t *__restrict__ action, UmmaWork work, uint32_t seed, uint32_t draw, int do_sample,
                   float *__restrict__ out_logits, int32_t *__restrict__ out_index)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~(uintptr_t)1023u);
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t w_base = smem_base, a_base = smem_base + kURing * kUBlockBytes, bar_base = a_base + 2 * kUBlockBytes;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + (kURing + 2) * kUBlockBytes + 128);
    UmmaSched *sched = reinterpret_cast<UmmaSched *>(smem + (kURing + 2) * kUBlockBytes + 160);
    static_assert(160 + sizeof(UmmaSched) <= 256, "the schedule sits between the barriers and the biases");
    /* barriers: 8 bytes each */
    auto W_FULL = [&](int i) { return bar_base + (uint32_t)i * 8u; };
    auto W_EMPTY = [&](int i) { return bar_base + 24u + (uint32_t)i * 8u; };
    auto ACC_FULL = [&](int t) { return bar_base + 48u + (uint32_t)t * 8u; };
    auto A_READY = [&](int t) { return bar_base + 64u + (uint32_t)t * 8u; };

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < kURing; ++i) { mbar_init(W_FULL(i), 1); mbar_init(W_EMPTY(i), 1); }
        for (int t = 0; t < 2; ++t) { mbar_init(ACC_FULL(t), 1); mbar_init(A_READY(t), kUEpiWarps / 2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        /* the schedule: rows of every network (from the launch, or from the totals the last row scan wrote -- the host then
         * neither knows nor waits for the population) and the tile pairs they make, numbered through all networks */
        uint32_t first = 0, pairs = 0;
        if (work.totals != nullptr)
            for (int k = 0; k < work.species0; ++k) first += work.totals->speciesTotal[k];
        for (int k = 0; k < kUMaxNets; ++k) {
            uint32_t rb = 0, rows = 0;
            if (k < work.n) {
                rb = work.totals != nullptr ? first : work.net[k].row_begin;
                rows = work.totals != nullptr ? work.totals->speciesTotal[work.species0 + k] : work.net[k].rows;
                first += rows;
            }
            pairs += ((rows + kUTile - 1) / kUTile + 1) / 2;
            sched->row_begin[k] = rb;
            sched->rows[k] = rows;
            sched->pair_end[k] = pairs;
            sched->out_base[k] = work.totals != nullptr ? rb : 0u;   /* device-resolved rows: outputs indexed by observation row */
        }
        sched->pairs = pairs;
    }
    if (warp == kUEpiWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"((uint32_t)__cvta_generic_to_shared(tmem_slot)), "r"((uint32_t)kUTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    float *bias_s = reinterpret_cast<float *>(smem + (kURing + 2) * kUBlockBytes + 256);
    const uint32_t bias_base = bar_base + 256u;
    for (int k = 0; k < work.n; ++k)
        for (int i = 0; i < work.net[k].n_ops; ++i) {
            const float *src = reinterpret_cast<const float *>(work.net[k].blob + work.net[k].b_ofs[i]);
            for (int j = tid; j < work.net[k].bias_n[i]; j += kUThreads) bias_s[work.net[k].bias_at[i] + j] = __ldg(src + j);
        }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    /* TMEM columns: tile t accumulator at t * 256, its scratch (LSTM cell / GRU helpers) at t * 256 + 128 */
    const uint32_t n_pairs = sched->pairs;

    if (warp < kUEpiWarps) {
        /* ---------------- epilogue warps: thread = one agent row of tile t, half of its columns ---------------- */
        const int t = warp >> 3, w4 = warp & 3, hf = (warp >> 2) & 1;
        const uint32_t r = (uint32_t)(w4 * 32 + lane);
        const uint32_t a_tile = a_base + (uint32_t)t * kUBlockBytes;        /* shared-space address of this tile's A operand */
        const uint32_t acc = tmem + ((uint32_t)(w4 * 32) << 16) + (uint32_t)t * 256u, scratch = acc + 128u;
        const Key draw_key = key_split(key_init(seed), draw);
        uint32_t acc_phase = 0;
        int k = 0;                                                           /* network of the current pair */
        UTRACE_DECL(1 + t * 2 + hf)
        for (uint32_t p = blockIdx.x; p < n_pairs; p += gridDim.x) {
            while (p >= sched->pair_end[k]) ++k;
            const UmmaNet &net = work.net[k];
            const uint32_t tile = (p - (k > 0 ? sched->pair_end[k - 1] : 0u)) * 2u + (uint32_t)t;
            const size_t lr = (size_t)tile * kUTile + r;
            const bool live = lr < sched->rows[k];
            const uint32_t row_begin = sched->row_begin[k];
            /* construct_obs of this row -> A operand (80 columns = 10 chunks; the MMA reads 5 k-steps) */
            {
                const size_t row = (size_t)row_begin + (live ? lr : 0);
                float x[80];
#pragma unroll
                for (int i = 0; i < 80; ++i) x[i] = 0.f;
                if (live) {
                    const uint4 d0 = __ldg(reinterpret_cast<const uint4 *>(o.depth + row * 32)), d1 = __ldg(reinterpret_cast<const uint4 *>(o.depth + row * 32) + 1);
                    const uint4 s0 = __ldg(reinterpret_cast<const uint4 *>(o.semantic + row * 32)), s1 = __ldg(reinterpret_cast<const uint4 *>(o.semantic + row * 32) + 1);
                    const float hl = __ldg(o.health + row);
                    const float2 ps = __ldg(reinterpret_cast<const float2 *>(o.position) + row);
                    const float2 su = __ldg(reinterpret_cast<const float2 *>(o.surrounding) + row);
                    const uint32_t dw[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
                    const uint32_t sw[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        x[i] = (float)((dw[i >> 2] >> (8 * (i & 3))) & 0xffu);
                        x[35 + i] = (float)(int8_t)((sw[i >> 2] >> (8 * (i & 3))) & 0xffu);
                    }
                    x[32] = hl; x[33] = ps.x; x[34] = ps.y; x[67] = su.x; x[68] = su.y;
                }
#pragma unroll
                for (int j = 0; j < 10; ++j) {
                    const uint4 pk = make_uint4(upack_bf16(x[8 * j], x[8 * j + 1]), upack_bf16(x[8 * j + 2], x[8 * j + 3]),
                                                upack_bf16(x[8 * j + 4], x[8 * j + 5]), upack_bf16(x[8 * j + 6], x[8 * j + 7]));
                    if ((j >= 5) == (hf != 0)) sts128(a_tile + sw128_chunk_offset(r, (uint32_t)j), pk);
                }
            }
            fence_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(A_READY(t));

            const int B = net.n_blocks;
            for (int b = 0; b < B; ++b) {
                const int opi = net.op_of[b], opc = net.opcode[opi], gate = net.gate_of[b];
                const uint32_t bias = bias_base + 4u * net.bias_at[opi];
                mbar_wait(ACC_FULL(t), acc_phase);
                acc_phase ^= 1u;
                tc_fence_after();
                if (lane == 0 && w4 == 0) { UTRACE(1, t, b) }
                const int cb = hf * 64;
                if (opc == MBOTS_OP_LINEAR) {
                    switch (net.act[opi]) {
                    case MBOTS_ACT_TANH: epi_block<EM_LINEAR, MBOTS_ACT_TANH>(acc, scratch, bias, bias, a_tile, r, cb); break;
                    case MBOTS_ACT_ELU: epi_block<EM_LINEAR, MBOTS_ACT_ELU>(acc, scratch, bias, bias, a_tile, r, cb); break;
                    case MBOTS_ACT_LOGSIGMOID: epi_block<EM_LINEAR, MBOTS_ACT_LOGSIGMOID>(acc, scratch, bias, bias, a_tile, r, cb); break;
                    case MBOTS_ACT_LEAKYRELU: epi_block<EM_LINEAR, MBOTS_ACT_LEAKYRELU>(acc, scratch, bias, bias, a_tile, r, cb); break;
                    case MBOTS_ACT_RELU: epi_block<EM_LINEAR, MBOTS_ACT_RELU>(acc, scratch, bias, bias, a_tile, r, cb); break;
                    case MBOTS_ACT_SIGMOID: epi_block<EM_LINEAR, MBOTS_ACT_SIGMOID>(acc, scratch, bias, bias, a_tile, r, cb); break;
                    default: epi_block<EM_LINEAR, MBOTS_ACT_NONE>(acc, scratch, bias, bias, a_tile, r, cb); break;
                    }
                } else if (opc == MBOTS_OP_LSTM0) {                   /* biases: b_i, b_g, b_o */
                    const uint32_t bg = bias + 512u * (uint32_t)gate;
                    if (gate == 0) epi_block<EM_LSTM_I, 0>(acc, scratch, bg, bg, a_tile, r, cb);
                    else if (gate == 1) epi_block<EM_LSTM_G, 0>(acc, scratch, bg, bg, a_tile, r, cb);
                    else epi_block<EM_LSTM_O, 0>(acc, scratch, bg, bg, a_tile, r, cb);
                } else if (opc == MBOTS_OP_GRU0) {                    /* biases: b_r, b_z, b_in, b_hn */
                    const uint32_t bg = bias + 512u * (uint32_t)gate;
                    if (gate == 0) epi_block<EM_GRU_R, 0>(acc, scratch, bg, bias + 3u * 512u, a_tile, r, cb);
                    else if (gate == 1) epi_block<EM_GRU_Z, 0>(acc, scratch, bg, bg, a_tile, r, cb);
                    else epi_block<EM_GRU_N, 0>(acc, scratch, bg, bg, a_tile, r, cb);
                } else {                                              /* HEAD: 6 logits in columns 0..5 */
                    uint32_t v[16];
                    if (hf == 0) { tmem_ld16(acc, v); tmem_wait_ld(v); }
                    if (live && hf == 0) {
                        const size_t orow = (size_t)sched->out_base[k] + lr;
                        float l[6];
#pragma unroll
                        for (int q = 0; q < 6; ++q) l[q] = __uint_as_float(v[q]) + bias_s[net.bias_at[opi] + q];
                        if (out_logits) {
#pragma unroll
                            for (int q = 0; q < 6; ++q) out_logits[orow * 6 + q] = l[q];
                        }
                        if (do_sample) {
                            float m = l[0];
#pragma unroll
                            for (int q = 1; q < 6; ++q) m = fmaxf(m, l[q]);
                            float wgt[6];
#pragma unroll
                            for (int q = 0; q < 6; ++q) wgt[q] = __expf(l[q] - m);
                            const uint32_t srow = row_begin + (uint32_t)lr;
                            const int pick = sample_and_write(wgt, draw_key, agent_id(o.owner, o.worldOff, srow), srow, action);
                            if (out_index) out_index[orow] = pick;
                        }
                    }
                }
                if (b + 1 < B) {          /* accumulator drained, A operand (re)written: the next block may be issued */
                    if (lane == 0 && w4 == 0) { UTRACE(3, t, b) }
                    fence_async_smem();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0 && w4 == 0) { UTRACE(5, t, b) }
                    if (lane == 0) mbar_arrive(A_READY(t));
                }
            }
        }
    } else {
        /* ---------------- issue warp: weight ring + tcgen05.mma. Every lane runs the (warp-uniform) control flow,
         * one elected lane issues the copies, the products and the commits ---------------- */
        /* next block to fetch: pair, its network, index in that network's block list, ring slot */
        uint32_t f_p = blockIdx.x, f_b = 0, f_slot = 0;
        int f_k = 0;
        auto fetch = [&]() {
            if (f_p < n_pairs) {
                while (f_p >= sched->pair_end[f_k]) ++f_k;
                if (elect_one()) {
                    mbar_expect_tx(W_FULL(f_slot), kUBlockBytes);
                    bulk_copy_g2s(w_base + f_slot * kUBlockBytes, work.net[f_k].blob + work.net[f_k].ofs[f_b], kUBlockBytes, W_FULL(f_slot));
                }
                if (++f_b == (uint32_t)work.net[f_k].n_blocks) { f_b = 0; f_p += gridDim.x; }
                if (++f_slot == (uint32_t)kURing) f_slot = 0;
            }
        };
        for (int i = 0; i < kURing; ++i) fetch();
        uint32_t a_phase = 0, slot = 0, w_phase = 0, prev_slot = 0, prev_phase = 0;
        bool first = true;
        int k = 0;
        UTRACE_DECL(0)
        for (uint32_t p = blockIdx.x; p < n_pairs; p += gridDim.x) {
            while (p >= sched->pair_end[k]) ++k;
            const int B = work.net[k].n_blocks;
            for (int b = 0; b < B; ++b) {
                const bool head = work.net[k].opcode[work.net[k].op_of[b]] == MBOTS_OP_HEAD;
                const uint32_t idesc = umma_idesc(head ? 16u : 128u);
                const uint64_t bdesc = umma_desc(w_base + slot * kUBlockBytes);
                if (!first) {
                    /* The products of the previous block are done about now (they were issued at the end of the last
                     * iteration): refill its slot at once, before this warp goes to sleep on the tiles' operands. A weight
                     * block takes ~5,500 cycles to land; refilling after this block's products instead left it one block
                     * period less (r4s: acting kernel -1.3 %). */
                    mbar_wait(W_EMPTY(prev_slot), prev_phase);
                    if (lane == 0) { UTRACE(13, 0, b) }
                    fetch();
                }
                mbar_wait(W_FULL(slot), w_phase);
                if (lane == 0) { UTRACE(10, 0, b) }
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    mbar_wait(A_READY(t), a_phase);     /* suspending wait: a busy poll here takes issue slots from four epilogue warps */
                    if (lane == 0) { UTRACE(11, t, b) }
                    tc_fence_after();
                    const uint64_t adesc = umma_desc(a_base + (uint32_t)t * kUBlockBytes);
                    if (elect_one()) {
                        if (b == 0) issue_mmas<5>(tmem + (uint32_t)t * 256u, adesc, bdesc, idesc);      /* the observation row: k_pad 80 */
                        else issue_mmas<8>(tmem + (uint32_t)t * 256u, adesc, bdesc, idesc);
                        umma_commit(ACC_FULL(t));
                    }
                    __syncwarp();
                    if (lane == 0) { UTRACE(12, t, b) }
                }
                a_phase ^= 1u;
                if (elect_one()) umma_commit(W_EMPTY(slot));
                __syncwarp();
                first = false;
                prev_slot = slot;
                prev_phase = w_phase;
                if (++slot == (uint32_t)kURing) { slot = 0; w_phase ^= 1u; }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kUEpiWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)kUTmemCols) : "memory");
}

/* Host: validate a SW128-layout program against the blob and derive the network's streaming plan; its biases go to
 * [*bias_floats, ...) of the shared-memory copy. Returns 0 or a message. */
static const char *policy_umma_plan(const mbots_policy_program &p, const void *blob, uint64_t blob_bytes, UmmaNet &net, uint32_t *bias_floats)
{
    if (p.n_ops < 2 || p.n_ops > MBOTS_POLICY_MAX_OPS) return "program must hold 2..12 ops";
    if (reinterpret_cast<uintptr_t>(blob) & 15u) return "the blob must be 16-byte aligned";
    net.blob = static_cast<const unsigned char *>(blob);
    net.n_blocks = 0;
    net.n_ops = p.n_ops;
    for (int i = 0; i < p.n_ops; ++i) {
        const mbots_policy_op &op = p.ops[i];
        const bool first = i == 0, last = i == p.n_ops - 1;
        if (first && (op.op != MBOTS_OP_LINEAR || op.k_pad != 80u)) return "the first op must be LINEAR with k_pad 80";
        if (!first && op.k_pad != 128u) return "every op after the first takes 128 inputs";
        if (last != (op.op == MBOTS_OP_HEAD)) return "HEAD must be the last op and only the last";
        if (op.op < MBOTS_OP_LINEAR || op.op > MBOTS_OP_HEAD) return "unknown op";
        if (op.act < MBOTS_ACT_NONE || op.act > MBOTS_ACT_SIGMOID) return "unknown activation";
        if ((op.w_ofs | op.b_ofs) & 15u) return "offsets must be multiples of 16";
        const uint32_t blocks = (op.op == MBOTS_OP_LSTM0 || op.op == MBOTS_OP_GRU0) ? 3u : 1u;
        const uint32_t nbias = op.op == MBOTS_OP_HEAD ? 8u : op.op == MBOTS_OP_GRU0 ? 4u * 128u : blocks * 128u;
        if ((uint64_t)op.w_ofs + (uint64_t)blocks * kUBlockBytes > blob_bytes || (uint64_t)op.b_ofs + 4ull * nbias > blob_bytes)
            return "op reaches beyond the blob";
        if (*bias_floats + nbias > (uint32_t)kUBiasFloats) return "the networks' biases exceed the kernel's shared-memory copy";
        net.opcode[i] = (uint8_t)op.op;
        net.act[i] = (uint8_t)op.act;
        net.b_ofs[i] = op.b_ofs;
        net.bias_at[i] = (uint16_t)*bias_floats;
        net.bias_n[i] = (uint16_t)nbias;
        *bias_floats += (nbias + 3u) & ~3u;               /* keeps every vector 16-byte aligned (ld.shared.v4) */
        for (uint32_t j = 0; j < blocks; ++j) {
            net.ofs[net.n_blocks] = op.w_ofs + j * (uint32_t)kUBlockBytes;
            net.op_of[net.n_blocks] = (uint8_t)i;
            net.gate_of[net.n_blocks] = (uint8_t)j;
            ++net.n_blocks;
        }
    }
    return nullptr;
}

static const char *launch_umma(const ObsSet &o, UmmaWork &work, uint32_t pairs_known, uint32_t seed, uint32_t draw, int do_sample,
                               float *out_logits, int32_t *out_index, cudaStream_t s)
{
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    static bool attr_set[64] = {};                      /* the opt-in is per device */
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        if (cudaFuncSetAttribute(policy_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kUSmemBytes) != cudaSuccess)
            return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed";
        attr_set[dev] = true;
    }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    /* persistent: one CTA per SM (all of them when the row counts are only known on the device) */
    const uint32_t grid = work.totals == nullptr && pairs_known < (uint32_t)sms ? pairs_known : (uint32_t)sms;
    const UObsPtrs ptrs = {o.depth, o.health, o.position, o.semantic, o.surrounding, o.owner, o.worldIdOffset};
    policy_umma_kernel<<<grid, kUThreads, kUSmemBytes, s>>>(ptrs, o.action, work, seed, draw, do_sample, out_logits, out_index);
    return nullptr;
}

const char *launch_policy_umma(const ObsSet &o, uint32_t row_begin, uint32_t rows, const void *blob, uint64_t blob_bytes,
                               const mbots_policy_program &prog, uint32_t seed, uint32_t draw, int do_sample,
                               float *out_logits, int32_t *out_index, cudaStream_t s)
{
    UmmaWork work = {};
    uint32_t bias_floats = 0;
    const char *err = policy_umma_plan(prog, blob, blob_bytes, work.net[0], &bias_floats);
    if (err || rows == 0) return err;
    work.n = 1;
    work.species0 = -1;
    work.totals = nullptr;
    work.net[0].row_begin = row_begin;
    work.net[0].rows = rows;
    return launch_umma(o, work, ((rows + kUTile - 1) / kUTile + 1) / 2, seed, draw, do_sample, out_logits, out_index, s);
}

/* `n` networks for the blocks of species species0 .. species0 + n - 1 of the current row set, the row ranges resolved on the
 * device from `totals`: ONE launch (n = 4: every agent of the simulator acts) */
const char *launch_policy_umma_species(const ObsSet &o, const Totals *totals, int species0, int n, const void *const *blobs,
                                       const uint64_t *blob_bytes, const mbots_policy_program *const *progs, uint32_t seed,
                                       uint32_t draw, int do_sample, float *out_logits, int32_t *out_index, cudaStream_t s)
{
    if (n < 1 || species0 < 0 || species0 + n > kUMaxNets) return "species range outside 0..3";
    UmmaWork work = {};
    uint32_t bias_floats = 0;
    for (int k = 0; k < n; ++k) {
        const char *err = policy_umma_plan(*progs[k], blobs[k], blob_bytes[k], work.net[k], &bias_floats);
        if (err) return err;
    }
    work.n = n;
    work.species0 = species0;
    work.totals = totals;
    return launch_umma(o, work, 0, seed, draw, do_sample, out_logits, out_index, s);
}

} // namespace mbk

#ifdef MBK_UMMA_TRACE
extern "C" int mbots_debug_umma_trace(unsigned long long *out, int reset)
{
    cudaDeviceSynchronize();
    if (out) cudaMemcpyFromSymbol(out, mbk::g_umma_trace, sizeof(mbk::g_umma_trace));
    if (reset) { void *p = nullptr; cudaGetSymbolAddress(&p, mbk::g_umma_trace); cudaMemset(p, 0, sizeof(mbk::g_umma_trace)); }
    return 0;
}
#endif
