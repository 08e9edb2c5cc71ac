/* policy_fused.cu -- one kernel per species for the whole acting path of learn/training_loop.py:52-83,136:
 * construct_obs -> feature MLP -> one recurrent step -> actor head -> softmax -> Categorical sample -> one-hot
 * write into the action export. Activations never leave the SM: a CTA owns a tile of 128 agents (8 warps x 16
 * rows), every layer is a [16 x K] x [K x 128] tensor-core product per warp (mma.sync m16n8k16, bf16 inputs,
 * fp32 accumulation), and a layer's accumulator fragments are re-packed in registers as the next layer's A
 * fragments. Weights (bf16, pre-padded on the host to the shared-memory layout) are streamed block by block
 * from L2 into a double buffer with cp.async while the previous block is being multiplied.
 *
 * The network is described by a short program (PolicyProgram) built by madrona_bots/policy.py from a
 * reference-format checkpoint (learn/ckpt.py:28-32, learn/models.py:95-165): LINEAR blocks with an activation,
 * the recurrent layer as one step from a zero state (RNN / LSTM / GRU; see policy.py for that position), HEAD.
 */
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "sim_dev.h"
#include "policy_dev.cuh"
#include "../../include/mbots_policy.h"

namespace mbk {

constexpr int kTileRows = 128;           /* agents per CTA tile */
constexpr int kHid = 128;                /* hidden width of every block */
constexpr int kPadK = 8;                 /* bf16 elements of row padding: 16 B, keeps ldmatrix conflict-free */
constexpr int kObsK = 80;                /* 69 observation columns padded to a multiple of 16 */
constexpr int kWBufBytes = kHid * (kHid + kPadK) * 2;      /* largest weight block: 34,816 B */
constexpr int kObsBytes = kTileRows * (kObsK + kPadK) * 2; /* 22,528 B */
constexpr int kFusedThreads = 256;

constexpr int kWBufs = 3;                /* weight blocks in flight: the one in use + two being fetched */
size_t policy_fused_smem_bytes() { return kWBufs * (size_t)kWBufBytes + kObsBytes; }

/* ---- tensor-core and async-copy primitives ---- */
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t smem_addr)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr));
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void *gptr)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi)
{
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
}

/* ---- activations (learn/models.py:63-69 + the gates of the recurrent layers). Hardware approximations
 * (tanh.approx, ex2/lg2 based exp/log: ~1e-3 absolute) -- far inside the bf16 rounding of every layer input. ---- */
__device__ __forceinline__ float tanh_fast(float x)
{
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sigmoidf_(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }

__device__ __forceinline__ float activate(int act, float x)
{
    switch (act) {
    case MBOTS_ACT_TANH: return tanh_fast(x);
    case MBOTS_ACT_ELU: return x > 0.f ? x : __expf(x) - 1.f;
    case MBOTS_ACT_LOGSIGMOID: return fminf(x, 0.f) - __logf(1.f + __expf(-fabsf(x)));
    case MBOTS_ACT_LEAKYRELU: return x > 0.f ? x : 0.01f * x;
    case MBOTS_ACT_RELU: return fmaxf(x, 0.f);
    case MBOTS_ACT_SIGMOID: return sigmoidf_(x);
    default: return x;
    }
}

/* ---- per-warp [16 x K] x [K x 128] product: A fragments in registers, B by ldmatrix from the weight block ---- */
template <int KS>
__device__ __forceinline__ void gemm_block(float (&acc)[16][4], const uint32_t (&a)[8][4], uint32_t wbase, uint32_t stride)
{
    const int lane = threadIdx.x & 31;
    /* ldmatrix.x4 = n-tile pair x {k lo, k hi}: lane l addresses row (l & 7) of matrix (l >> 3) */
    const uint32_t lane_addr = wbase + (uint32_t)(((lane >> 4) << 3) + (lane & 7)) * stride + (uint32_t)((lane >> 3) & 1) * 16u;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) { acc[nt][0] = 0.f; acc[nt][1] = 0.f; acc[nt][2] = 0.f; acc[nt][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
        for (int np = 0; np < 8; ++np) {
            uint32_t b[4];
            ldmatrix_x4(b, lane_addr + (uint32_t)(np * 16) * stride + (uint32_t)ks * 32u);
            mma_bf16(acc[2 * np], a[ks], b[0], b[1]);
            mma_bf16(acc[2 * np + 1], a[ks], b[2], b[3]);
        }
    }
}

/* accumulator fragments (+ bias, activation) -> A fragments of the next block */
__device__ __forceinline__ void repack(uint32_t (&a)[8][4], const float (&v)[16][4])
{
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
        a[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(v[nt][0], v[nt][1]);     /* row g */
        a[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(v[nt][2], v[nt][3]);     /* row g + 8 */
    }
}

template <int ACT>
__device__ __forceinline__ void linear_epilogue(float (&acc)[16][4], const unsigned char *__restrict__ bias, int q)
{
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
        const float2 bb = __ldg(reinterpret_cast<const float2 *>(bias) + nt * 4 + q);
        acc[nt][0] = activate(ACT, acc[nt][0] + bb.x); acc[nt][1] = activate(ACT, acc[nt][1] + bb.y);
        acc[nt][2] = activate(ACT, acc[nt][2] + bb.x); acc[nt][3] = activate(ACT, acc[nt][3] + bb.y);
    }
}

struct FusedPlan {                        /* weight blocks in streaming order */
    int32_t n_blocks;
    uint32_t ofs[3 * MBOTS_POLICY_MAX_OPS];
    uint32_t bytes[3 * MBOTS_POLICY_MAX_OPS];
    uint8_t op_of[3 * MBOTS_POLICY_MAX_OPS];      /* op a block belongs to */
    uint8_t gate_of[3 * MBOTS_POLICY_MAX_OPS];    /* its index within that op (recurrent gates) */
};

struct ObsPtrs { const uint8_t *depth; const float *health; const float *position; const int8_t *semantic; const float *surrounding;
                 const uint32_t *owner; uint32_t worldOff; };

__global__ void __launch_bounds__(kFusedThreads, 1)
policy_fused_kernel(ObsPtrs o, int32_t *__restrict__ action, uint32_t row_begin, uint32_t rows, const unsigned char *__restrict__ blob,
                    mbots_policy_program prog, FusedPlan plan, uint32_t seed, uint32_t draw, int do_sample,
                    float *__restrict__ out_logits, int32_t *__restrict__ out_index)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t obs_base = smem_base + (uint32_t)kWBufs * kWBufBytes;
    __nv_bfloat16 *obs_tile = reinterpret_cast<__nv_bfloat16 *>(smem + kWBufs * kWBufBytes);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, q = lane & 3;
    const Key draw_key = key_split(key_init(seed), draw);

    auto issue_block = [&](int b) {       /* all threads: 16-byte chunks of block b into buffer b % kWBufs */
        if (b < plan.n_blocks) {
            const unsigned char *src = blob + plan.ofs[b];
            const uint32_t dst = smem_base + (uint32_t)(b % kWBufs) * kWBufBytes;
            for (uint32_t c = (uint32_t)tid * 16u; c < plan.bytes[b]; c += kFusedThreads * 16u) cp_async16(dst + c, src + c);
        }
        cp_async_commit();
    };
    /* block b has landed (only the group of block b + 1 may still be in flight) and every warp is done with block
     * b - 1, whose buffer the fetch of block b + 2 now reuses */
    auto acquire_block = [&](int b) -> uint32_t {
        cp_async_wait_but_one();
        __syncthreads();
        issue_block(b + 2);
        return smem_base + (uint32_t)(b % kWBufs) * kWBufBytes;
    };
    auto bias2 = [&](uint32_t b_ofs, int vec, int nt) -> float2 {
        return __ldg(reinterpret_cast<const float2 *>(blob + b_ofs + (size_t)vec * (kHid * 4)) + nt * 4 + q);
    };

    for (uint32_t tile = blockIdx.x; (size_t)tile * kTileRows < rows; tile += gridDim.x) {
        __syncthreads();                  /* the previous tile no longer reads shared memory */
        issue_block(0);
        issue_block(1);
        /* construct_obs of this tile's rows, bf16, zero padded to 80 columns / 128 rows: two threads per row, each
         * converts 16 depth and 16 semantic bytes (one 16-byte load each); the even one adds the five floats.
         * Every load of a thread is issued before the first conversion: one DRAM round trip per tile. */
        {
            const int r = tid >> 1, h = tid & 1;
            const size_t lr = (size_t)tile * kTileRows + r;
            const bool live = lr < rows;
            const size_t row = (size_t)row_begin + (live ? lr : 0);
            uint4 dv = make_uint4(0, 0, 0, 0), sv = make_uint4(0, 0, 0, 0);
            float hl = 0.f;
            float2 ps = make_float2(0.f, 0.f), su = make_float2(0.f, 0.f);
            if (live) {
                dv = __ldg(reinterpret_cast<const uint4 *>(o.depth + row * 32) + h);
                sv = __ldg(reinterpret_cast<const uint4 *>(o.semantic + row * 32) + h);
                if (h == 0) {
                    hl = __ldg(o.health + row);
                    ps = __ldg(reinterpret_cast<const float2 *>(o.position) + row);
                    su = __ldg(reinterpret_cast<const float2 *>(o.surrounding) + row);
                }
            }
            __nv_bfloat16 *dst = obs_tile + r * (kObsK + kPadK);
            const uint32_t dw[4] = {dv.x, dv.y, dv.z, dv.w}, sw[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                dst[h * 16 + i] = __float2bfloat16((float)((dw[i >> 2] >> (8 * (i & 3))) & 0xffu));
                dst[35 + h * 16 + i] = __float2bfloat16((float)(int8_t)((sw[i >> 2] >> (8 * (i & 3))) & 0xffu));
            }
            if (h == 0) {
                dst[32] = __float2bfloat16(hl);
                dst[33] = __float2bfloat16(ps.x); dst[34] = __float2bfloat16(ps.y);
                dst[67] = __float2bfloat16(su.x); dst[68] = __float2bfloat16(su.y);
            } else {
#pragma unroll
                for (int c = kObsDim; c < kObsK; ++c) dst[c] = __float2bfloat16(0.f);
            }
        }

        /* One pass per weight block: product, then the epilogue that block's (op, gate) asks for. Two product
         * instantiations in the whole kernel (K = 80 and K = 128) keep the code inside the instruction cache. */
        uint32_t a[8][4];
        float acc[16][4];
        float keep[16][4];      /* LSTM: sigmoid(i), then c. GRU: [0..1] = r * b_hn, [2..3] = 1 - z, as packed bf16 pairs */
#pragma unroll 1
        for (int b = 0; b < plan.n_blocks - 1; ++b) {
            const mbots_policy_op op = prog.ops[plan.op_of[b]];
            const int gate = plan.gate_of[b];
            const uint32_t w = acquire_block(b);
            if (b == 0) {
                /* first block: A fragments from the observation tile */
                const uint32_t row_addr = obs_base + (uint32_t)(warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * ((kObsK + kPadK) * 2) +
                                          (uint32_t)(lane >> 4) * 16u;
#pragma unroll
                for (int ks = 0; ks < kObsK / 16; ++ks) ldmatrix_x4(a[ks], row_addr + (uint32_t)ks * 32u);
                gemm_block<kObsK / 16>(acc, a, w, (kObsK + kPadK) * 2);
            } else {
                gemm_block<kHid / 16>(acc, a, w, (kHid + kPadK) * 2);
            }
            if (op.op == MBOTS_OP_LINEAR) {
                switch (op.act) {
                case MBOTS_ACT_TANH: linear_epilogue<MBOTS_ACT_TANH>(acc, blob + op.b_ofs, q); break;
                case MBOTS_ACT_ELU: linear_epilogue<MBOTS_ACT_ELU>(acc, blob + op.b_ofs, q); break;
                case MBOTS_ACT_LOGSIGMOID: linear_epilogue<MBOTS_ACT_LOGSIGMOID>(acc, blob + op.b_ofs, q); break;
                case MBOTS_ACT_LEAKYRELU: linear_epilogue<MBOTS_ACT_LEAKYRELU>(acc, blob + op.b_ofs, q); break;
                case MBOTS_ACT_RELU: linear_epilogue<MBOTS_ACT_RELU>(acc, blob + op.b_ofs, q); break;
                case MBOTS_ACT_SIGMOID: linear_epilogue<MBOTS_ACT_SIGMOID>(acc, blob + op.b_ofs, q); break;
                default: linear_epilogue<MBOTS_ACT_NONE>(acc, blob + op.b_ofs, q); break;
                }
                repack(a, acc);
            } else if (op.op == MBOTS_OP_LSTM0) {
                if (gate == 0) {                                      /* i gate */
#pragma unroll
                    for (int nt = 0; nt < 16; ++nt) {
                        const float2 bb = bias2(op.b_ofs, 0, nt);
                        keep[nt][0] = sigmoidf_(acc[nt][0] + bb.x); keep[nt][1] = sigmoidf_(acc[nt][1] + bb.y);
                        keep[nt][2] = sigmoidf_(acc[nt][2] + bb.x); keep[nt][3] = sigmoidf_(acc[nt][3] + bb.y);
                    }
                } else if (gate == 1) {                               /* g gate: c = sigmoid(i) * tanh(g) */
#pragma unroll
                    for (int nt = 0; nt < 16; ++nt) {
                        const float2 bb = bias2(op.b_ofs, 1, nt);
                        keep[nt][0] *= tanh_fast(acc[nt][0] + bb.x); keep[nt][1] *= tanh_fast(acc[nt][1] + bb.y);
                        keep[nt][2] *= tanh_fast(acc[nt][2] + bb.x); keep[nt][3] *= tanh_fast(acc[nt][3] + bb.y);
                    }
                } else {                                              /* o gate: h = sigmoid(o) * tanh(c) */
#pragma unroll
                    for (int nt = 0; nt < 16; ++nt) {
                        const float2 bb = bias2(op.b_ofs, 2, nt);
                        acc[nt][0] = sigmoidf_(acc[nt][0] + bb.x) * tanh_fast(keep[nt][0]); acc[nt][1] = sigmoidf_(acc[nt][1] + bb.y) * tanh_fast(keep[nt][1]);
                        acc[nt][2] = sigmoidf_(acc[nt][2] + bb.x) * tanh_fast(keep[nt][2]); acc[nt][3] = sigmoidf_(acc[nt][3] + bb.y) * tanh_fast(keep[nt][3]);
                    }
                    repack(a, acc);
                }
            } else {                                                  /* GRU0 */
                if (gate == 0) {                                      /* r gate */
#pragma unroll
                    for (int nt = 0; nt < 16; ++nt) {
                        const float2 bb = bias2(op.b_ofs, 0, nt), hn = bias2(op.b_ofs, 3, nt);
                        keep[nt][0] = __uint_as_float(pack_bf16(sigmoidf_(acc[nt][0] + bb.x) * hn.x, sigmoidf_(acc[nt][1] + bb.y) * hn.y));
                        keep[nt][1] = __uint_as_float(pack_bf16(sigmoidf_(acc[nt][2] + bb.x) * hn.x, sigmoidf_(acc[nt][3] + bb.y) * hn.y));
                    }
                } else if (gate == 1) {                               /* z gate */
#pragma unroll
                    for (int nt = 0; nt < 16; ++nt) {
                        const float2 bb = bias2(op.b_ofs, 1, nt);
                        keep[nt][2] = __uint_as_float(pack_bf16(1.f - sigmoidf_(acc[nt][0] + bb.x), 1.f - sigmoidf_(acc[nt][1] + bb.y)));
                        keep[nt][3] = __uint_as_float(pack_bf16(1.f - sigmoidf_(acc[nt][2] + bb.x), 1.f - sigmoidf_(acc[nt][3] + bb.y)));
                    }
                } else {                                              /* n gate: h = (1 - z) * tanh(n + b_in + r * b_hn) */
#pragma unroll
                    for (int nt = 0; nt < 16; ++nt) {
                        const float2 bb = bias2(op.b_ofs, 2, nt);
                        const float2 r0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&keep[nt][0]));
                        const float2 r1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&keep[nt][1]));
                        const float2 z0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&keep[nt][2]));
                        const float2 z1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&keep[nt][3]));
                        acc[nt][0] = z0.x * tanh_fast(acc[nt][0] + bb.x + r0.x); acc[nt][1] = z0.y * tanh_fast(acc[nt][1] + bb.y + r0.y);
                        acc[nt][2] = z1.x * tanh_fast(acc[nt][2] + bb.x + r1.x); acc[nt][3] = z1.y * tanh_fast(acc[nt][3] + bb.y + r1.y);
                    }
                    repack(a, acc);
                }
            }
        }
        {                                                             /* HEAD: 8 output rows, one n-tile */
            const int b = plan.n_blocks - 1;
            const mbots_policy_op op = prog.ops[prog.n_ops - 1];
            const uint32_t w = acquire_block(b);
            float lg[4] = {0.f, 0.f, 0.f, 0.f};
            const uint32_t lane_addr = w + (uint32_t)(lane & 7) * ((kHid + kPadK) * 2) + (uint32_t)(lane >> 3) * 16u;
#pragma unroll
            for (int kp = 0; kp < kHid / 32; ++kp) {                  /* two k-steps per ldmatrix.x4 */
                uint32_t bf[4];
                ldmatrix_x4(bf, lane_addr + (uint32_t)kp * 64u);
                mma_bf16(lg, a[2 * kp], bf[0], bf[1]);
                mma_bf16(lg, a[2 * kp + 1], bf[2], bf[3]);
            }
            const float2 bb = __ldg(reinterpret_cast<const float2 *>(blob + op.b_ofs) + q);
            lg[0] += bb.x; lg[1] += bb.y; lg[2] += bb.x; lg[3] += bb.y;
            /* the 8 logits of a row sit in the 4 lanes of a quad: lane q = 0 takes row g, lane q = 1 row g + 8 */
            float l0[8], l1[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int src = (lane & ~3) + j;
                l0[2 * j] = __shfl_sync(0xffffffffu, lg[0], src); l0[2 * j + 1] = __shfl_sync(0xffffffffu, lg[1], src);
                l1[2 * j] = __shfl_sync(0xffffffffu, lg[2], src); l1[2 * j + 1] = __shfl_sync(0xffffffffu, lg[3], src);
            }
            if (q < 2) {
                const size_t lr = (size_t)tile * kTileRows + warp * 16 + g + (q ? 8 : 0);
                if (lr < rows) {
                    float l[6];
#pragma unroll
                    for (int k = 0; k < 6; ++k) l[k] = q ? l1[k] : l0[k];
                    if (out_logits) {
#pragma unroll
                        for (int k = 0; k < 6; ++k) out_logits[lr * 6 + k] = l[k];
                    }
                    if (do_sample) {
                        float m = l[0];
#pragma unroll
                        for (int k = 1; k < 6; ++k) m = fmaxf(m, l[k]);
                        float wgt[6];
#pragma unroll
                        for (int k = 0; k < 6; ++k) wgt[k] = __expf(l[k] - m);
                        const uint32_t srow = row_begin + (uint32_t)lr;
                        const int pick = sample_and_write(wgt, draw_key, agent_id(o.owner, o.worldOff, srow), srow, action);
                        if (out_index) out_index[lr] = pick;
                    }
                }
            }
        }
    }
    cp_async_wait_all();
}

/* Host: validate the program against the blob size and derive the streaming plan. Returns 0 or a message. */
const char *policy_plan(const mbots_policy_program &p, uint64_t blob_bytes, FusedPlan &plan)
{
    if (p.n_ops < 2 || p.n_ops > MBOTS_POLICY_MAX_OPS) return "program must hold 2..12 ops";
    plan.n_blocks = 0;
    for (int i = 0; i < p.n_ops; ++i) {
        const mbots_policy_op &op = p.ops[i];
        const bool first = i == 0, last = i == p.n_ops - 1;
        if (first && (op.op != MBOTS_OP_LINEAR || op.k_pad != (uint32_t)kObsK)) return "the first op must be LINEAR with k_pad 80";
        if (!first && op.k_pad != (uint32_t)kHid) return "every op after the first takes 128 inputs";
        if (last != (op.op == MBOTS_OP_HEAD)) return "HEAD must be the last op and only the last";
        if (op.op < MBOTS_OP_LINEAR || op.op > MBOTS_OP_HEAD) return "unknown op";
        if (op.act < MBOTS_ACT_NONE || op.act > MBOTS_ACT_SIGMOID) return "unknown activation";
        if ((op.w_ofs | op.b_ofs) & 15u) return "offsets must be multiples of 16";
        const uint32_t rows = op.op == MBOTS_OP_HEAD ? 8u : (uint32_t)kHid;
        const uint32_t blocks = (op.op == MBOTS_OP_LSTM0 || op.op == MBOTS_OP_GRU0) ? 3u : 1u;
        const uint32_t bytes = rows * (op.k_pad + kPadK) * 2u;
        const uint32_t nbias = op.op == MBOTS_OP_HEAD ? 8u : op.op == MBOTS_OP_GRU0 ? 4u * kHid : blocks * kHid;
        if ((uint64_t)op.w_ofs + (uint64_t)blocks * bytes > blob_bytes || (uint64_t)op.b_ofs + 4ull * nbias > blob_bytes)
            return "op reaches beyond the blob";
        for (uint32_t j = 0; j < blocks; ++j) {
            plan.ofs[plan.n_blocks] = op.w_ofs + j * bytes;
            plan.bytes[plan.n_blocks] = bytes;
            plan.op_of[plan.n_blocks] = (uint8_t)i;
            plan.gate_of[plan.n_blocks] = (uint8_t)j;
            ++plan.n_blocks;
        }
    }
    return nullptr;
}

const char *launch_policy_fused(const ObsSet &o, uint32_t row_begin, uint32_t rows, const void *blob, uint64_t blob_bytes,
                                const mbots_policy_program &prog, uint32_t seed, uint32_t draw, int do_sample,
                                float *out_logits, int32_t *out_index, cudaStream_t s)
{
    FusedPlan plan;
    const char *err = policy_plan(prog, blob_bytes, plan);
    if (err || rows == 0) return err;
    const size_t smem = policy_fused_smem_bytes();
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    static bool attr_set[64] = {};                      /* the opt-in is per device */
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        if (cudaFuncSetAttribute(policy_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed";
        attr_set[dev] = true;
    }
    const uint32_t tiles = (rows + kTileRows - 1) / kTileRows;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint32_t grid = tiles < (uint32_t)sms ? tiles : (uint32_t)sms;      /* persistent: one CTA per SM */
    const ObsPtrs ptrs = {o.depth, o.health, o.position, o.semantic, o.surrounding, o.owner, o.worldIdOffset};
    policy_fused_kernel<<<grid, kFusedThreads, smem, s>>>(ptrs, o.action, row_begin, rows,
                                                          static_cast<const unsigned char *>(blob), prog, plan, seed, draw,
                                                          do_sample, out_logits, out_index);
    return nullptr;
}

} // namespace mbk
