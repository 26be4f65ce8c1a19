// Grouped GEMM launch interface shared by i64b_gemm.cu (kernels) and i64b_api.cu (the block).
#pragma once
#include <cuda.h>
#include "i64b_common.cuh"

// Operand kinds: bf16 x bf16 -> fp32 (kind::f16), int8 x int8 -> int32 (kind::i8), E4M3 x E4M3 -> fp32 (kind::f8f6f4)
enum { I64B_TYPE_BF16 = 0, I64B_TYPE_INT8 = 1, I64B_TYPE_FP8 = 2 };

// Epilogue selection (runtime, within a template instantiation)
enum {
    I64B_EPI_F32 = 0,       // plain: out = acc (fp32 or bf16 store)
    I64B_EPI_DEQUANT = 1,   // int8: out = float(acc) * (row_scale * col_scale)
};

struct I64bGemmParams {
    const I64bGroup* groups;   // device group table, or nullptr → use `inline_group`
    I64bGroup inline_group;
    int32_t num_groups;
    int32_t k[2];              // reduction length per `which`
    void* out[2];              // output buffer per `which`
    int64_t ld_out[2];         // leading dimension of out (elements)
    int32_t out_dtype;         // I64B_BF16 | I64B_F32 (plain epilogues); SwiGLU-bf16 always bf16; Q14 always int32
    int32_t epi;               // I64B_EPI_*
    const float* row_scale[2]; // int8: per A-row scale, indexed by A row of buffer `which`
    const float* col_scale[2]; // int8: per weight-row (out-feature) scale of packed tensor `which`
    int32_t* raw_out;          // optional raw int32 accumulator dump (plain int8 epilogue, which==0 layout)
    int64_t ld_raw;
    int32_t* amax_part[2];     // Q14 epilogue: per (row, n_tile) max|q14|
    int32_t amax_ld[2];
    const int16_t* silu_lut;   // device copy of the Q7 SiLU table (2049 entries)
    // fused combine (GEMM2 only): the shared-expert tiles (last group, token order) pull the routed rows of their
    // tokens and write the finished block output  rs*(sum_k w_k*y_k) + ss*y_shared  (token_routed_mlp.py:156-185)
    int32_t fuse_combine;
    int32_t keep_shared_y;     // debug taps: also store the shared expert's own rows to out[1]
    const int32_t* dest_row;   // [K, N] row of assignment (k, t) in out[0]
    int32_t n_tokens, top_k;
    float w[I64B_MAX_TOPK];    // fused combine supports top_k <= 2
    float routed_scale, shared_scale;
    void* final_out;           // [N, ld_final] bf16 or fp32
    int32_t final_dtype;
    int64_t ld_final;
    int32_t* tile_done;        // counter of drained routed (tile, epilogue-warp) pairs, zero at launch
    // token-major layout (I64bTriWs).  GEMM1 (dual, bf16): output rows are scattered — rowmap[0][r] = pair-order row |
    // (route << 30) of dispatched row r, rowmap[1][t] = pair-order row of token t — and land at column tri_col[route]
    // (tri_col[2] for the shared expert) of the [N, ld_out] buffer.  The fused down projection (i64b_launch_gemm_tri)
    // reads it back with K segments k[0], k[0], k[1] and writes row row_tok[r] of final_out.
    const int32_t* rowmap[2];
    int32_t tri_col[3];
    const int32_t* row_tok;    // pair-order row -> token | (bit 30: the token's route 1 sits in K segment 0)
    int32_t tri_scaled;        // one-accumulator form: GEMM1 multiplies the SwiGLU rows by tri_mul (powers of two: exact)
    float tri_mul[3];          // w_0 * rs, w_1 * rs (by ROUTE, bit 31 of rowmap), ss
    // GEMM1 without a dispatch pass (pair kernel, bf16): the A rows of the routed groups (which == 0) are gathered from the
    // caller's x — tmA0 is then a one-row-box map over x — through gather_idx[dispatched row] = token (indices beyond
    // gather_rows, i.e. the padding of the last m-tile, are not read; every index is clamped to [0, n_tokens))
    const int32_t* gather_idx;
    int32_t gather_rows;
    // bring-up instrumentation (env I64B_DEBUG_FLAGS / I64B_DEBUG_TIMELINE; never set in production)
    uint32_t debug_flags;      // bit0: skip MMAs (pure TMA streaming), bit1: skip epilogue stores; pair kernels, timing only
                               // (garbage results): bit7 / bit8 = no B / no A load on odd k-blocks (operand-ingest probe)
    long long* debug_timeline; // [grid, 8] clock64 offsets per CTA
};

// Tensor map for a row-major [rows, cols] matrix of 1-byte (int8) or 2-byte (bf16) elements whose
// tiles are (box_rows x 128 bytes), 128B-swizzled.
int i64b_encode_tmap(CUtensorMap* out, const void* base, bool is_int8, int64_t rows, int64_t cols,
                     int64_t pitch_bytes, int box_rows);

// Launch one grouped GEMM.  n_mma in {64,128,256}; dual = gate|up pair + fused SwiGLU epilogue.
// max_tiles is an upper bound on the number of (group, m, n) tiles, used only to size the grid.
// ctas = 2 runs CTA pairs (cta_group::2, M = 256 per MMA; n_mma must be 256): the B tensor maps must then have
// 128-row boxes (each CTA stages half of the B tile) and max_tiles counts (256-row, n) units.
// p.gather_idx != nullptr (bf16, dual, ctas == 2 only): see I64bGemmParams::gather_idx.
int i64b_launch_gemm(int type /* I64B_TYPE_* */, bool dual, int n_mma, int ctas, const CUtensorMap& tmA0, const CUtensorMap& tmA1,
                     const CUtensorMap& tmB0, const CUtensorMap& tmB1, const I64bGemmParams& p,
                     int64_t max_tiles, cudaStream_t stream);

// Fused down projection + combine over the token-major layout (top-2 routed experts + optional shared expert), CTA
// pairs, 256 x 256 tiles, two TMEM accumulators (the routed sum is folded back into the first one):
//   out[row_tok[r]] = rs * (wa * A[r, 0:I] · W2[ea]^T + wb * A[r, I:2I] · W2[eb]^T) + ss * A[r, 2I:2I+S] · W2s^T
// in the reference's fp32 operation order (layers/token_routed_mlp.py:156-185).  tmA: [N, 2I+S] (128-row boxes);
// tmB0 / tmB1: routed / shared down-projection weights with 128-row boxes; p.groups: one group per UNORDERED expert pair
// ea <= eb (w_row_base = ea*H, pad0 = eb*H); (wa, wb) = (w0, w1), swapped for the rows whose route 0 is eb; p.k[0] = I, p.k[1] = S (0: no shared expert).  max_units: upper bound on the number of
// 256-row units.
// one_acc: the scales are already in the rows of A (GEMM1 epilogue, tri_mul — every factor a power of two), the three
// segments accumulate into ONE TMEM accumulator and the two accumulators alternate between tiles.
int i64b_launch_gemm_tri(int type /* I64B_TYPE_* */, bool one_acc, const CUtensorMap& tmA, const CUtensorMap& tmB0,
                         const CUtensorMap& tmB1, const I64bGemmParams& p, int64_t max_units, cudaStream_t stream);

int i64b_num_sms();
int i64b_set_gemm_cta_limit(int limit);      // thread-local; returns the previous value
void i64b_count_launch(int n = 1);
