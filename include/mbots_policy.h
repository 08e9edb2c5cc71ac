/* mbots_policy.h -- network description consumed by mbots_policy_act (include/mbots.h): the species
 * networks of the reference (learn/models.py:14-93: feature MLP, one recurrent layer, actor head; checkpoint
 * format learn/ckpt.py:28-32) flattened into a short program over one device blob.
 *
 * Blob layout (built by madrona_bots/policy.py, all offsets multiples of 16 bytes), two weight-block formats:
 *   MBOTS_LAYOUT_ROWPAD  weight block [rows][k_pad + 8] bfloat16, row = output feature, k = input feature, zero
 *                 padded; rows = 128 (HEAD: 8). The 8 trailing elements of a row are padding (shared-memory
 *                 layout of the mma.sync kernel, policy_fused.cu).
 *   MBOTS_LAYOUT_SW128   every weight block is 32 KB: the canonical K-major SWIZZLE_128B shared-memory image of a
 *                 [128 rows][128 k] bfloat16 operand that tcgen05.mma reads (policy_umma.cu): element (row, k) at
 *                 byte (k/64)*16384 + (row/8)*1024 + (row%8)*128 + (((k%64)/8) ^ (row%8))*16 + (k%8)*2; unused
 *                 rows / columns zero (HEAD: rows 0..5 used).
 *   bias vector   float32 (both layouts).
 * Ops, applied in order to a tile of 128 agents whose first input is the (69 -> k_pad 80) observation row of
 * learn/util.py:23-29:
 *   LINEAR  x <- act(W x + b)                       1 block, 128 biases
 *   LSTM0   one LSTM step from a zero state         3 blocks (i, g, o gates; the f gate multiplies c0 = 0),
 *           h = sigmoid(o) * tanh(sigmoid(i) * tanh(g))   3 x 128 biases (b_ih + b_hh of each gate)
 *   GRU0    one GRU step from a zero state          3 blocks (r, z, n gates),
 *           h = (1 - sigmoid(z)) * tanh(n + r * b_hn)     4 x 128 biases: b_r, b_z (ih + hh), b_in, b_hn
 *   HEAD    logits <- W x + b (6 of 8 rows used)    1 block of 8 rows, 8 biases; must be the last op
 */
#ifndef MBOTS_POLICY_H
#define MBOTS_POLICY_H
#include <stdint.h>

enum mbots_policy_opcode { MBOTS_OP_LINEAR = 0, MBOTS_OP_LSTM0 = 1, MBOTS_OP_GRU0 = 2, MBOTS_OP_HEAD = 3 };
enum mbots_policy_act {
    MBOTS_ACT_NONE = 0, MBOTS_ACT_TANH = 1, MBOTS_ACT_ELU = 2, MBOTS_ACT_LOGSIGMOID = 3,
    MBOTS_ACT_LEAKYRELU = 4, MBOTS_ACT_RELU = 5, MBOTS_ACT_SIGMOID = 6
};
enum mbots_policy_layout { MBOTS_LAYOUT_ROWPAD = 0, MBOTS_LAYOUT_SW128 = 1 };
#define MBOTS_POLICY_MAX_OPS 12

typedef struct mbots_policy_op {
    int32_t op;        /* mbots_policy_opcode */
    int32_t act;       /* mbots_policy_act (LINEAR only) */
    uint32_t k_pad;    /* input width padded to a multiple of 16: 80 for the first op, 128 after */
    uint32_t w_ofs;    /* byte offset of the first weight block in the blob */
    uint32_t b_ofs;    /* byte offset of the bias vector(s) in the blob */
} mbots_policy_op;

typedef struct mbots_policy_program {
    int32_t n_ops;
    int32_t layout;    /* mbots_policy_layout: which kernel the weight blocks were laid out for */
    mbots_policy_op ops[MBOTS_POLICY_MAX_OPS];
} mbots_policy_program;

#endif /* MBOTS_POLICY_H */
