// Integer / bandwidth-bound stages of the block: routing, deterministic dispatch plan,
// fused (quantize-and-)scatter, re-quantisation of the integer SwiGLU output, weighted combine.
#pragma once
#include "i64b_common.cuh"

enum { I64B_ROUTE_TABLE = 0, I64B_ROUTE_MODULO = 1, I64B_ROUTE_GIVEN = 2, I64B_ROUTE_NOIDS = 3 };

// Layout of the plan workspace (all int32 unless noted); sizes depend on (n, K, E) only.
struct I64bPlanWs {
    uint32_t* packed;      // [K, N]   expert (8 bits) | rank-in-tile (24 bits)
    int32_t* tile_counts;  // [tiles, K*E]   per-tile histogram, then in place: exclusive base over tiles
    int32_t* counts;       // [K, E]
    int32_t* offsets;      // [K, E]   per-route exclusive scan (reference convention)
    int32_t* route_base;   // [K, E]   rows of routes < k inside expert e's merged segment
    int32_t* group_start;  // [E + 1]  merged expert-major segment starts (last = N*K)
    I64bGroup* groups1;    // [E + 1]  GEMM1 problems (gate|up), shared expert last
    I64bGroup* groups2;    // [E + 1]  GEMM2 problems (down)
    int32_t* sync;         // [64]     device-side counters, zeroed by the plan ([0]: routed GEMM2 tiles drained)
    size_t bytes;
};
I64bPlanWs i64b_plan_ws_carve(void* base, int64_t n, int top_k, int num_experts);

struct I64bPlanArgs {
    int mode;                    // I64B_ROUTE_*
    const int64_t* token_ids;    // TABLE / MODULO
    const uint8_t* table;        // TABLE: [K, V] uint8
    const int32_t* given;        // GIVEN: [K, N] int32 expert ids
    int64_t n;
    int vocab, top_k, num_experts;
    // geometry for the group tables (0 when only the plan is wanted)
    int expert_inter, shared_inter, hidden;
    int* ticket;                 // optional persistent zero-initialised counter: the last rank CTA finalises the plan
                                 // in the same launch (self-resetting); nullptr -> separate one-CTA finalise launch
    int shared_from_x;           // 1: the shared expert's GEMM1 reads the caller's x (token order) — no dispatched copy
};

// route + per-tile stable ranks + finalisation (counts, offsets, group tables). 1-2 launches.
// tri != nullptr (top-2 blocks): the token-major layout's pair ranks and groups (I64bTriWs, below) in the same launch
int i64b_run_plan(const I64bPlanArgs& a, const I64bPlanWs& ws, cudaStream_t stream, const struct I64bTriWs* tri = nullptr);

// Token-major layout for the fused down projection (top-2 blocks): tokens stably sorted by their ordered expert pair
// (e0, e1).  All three SwiGLU intermediates of a token (route 0, route 1, shared expert) then sit in ONE row of a
// [N, 2*I_e + S] buffer, a pair's rows are contiguous, and a single grouped GEMM over K = 2*I_e + S with three TMEM
// accumulators per tile produces y0, y1 and y_shared of the same token in the same CTA: the combine happens in that
// GEMM's epilogue and the fp32 per-assignment outputs never exist in memory.
struct I64bTriWs {
    uint32_t* packed;     // [N]   pair id (8 bits) | rank inside the 1024-token tile (24 bits)
    int32_t* tile_base;   // [tiles, E*E]  per-tile histogram, then in place: exclusive base over tiles
    int32_t* pair_start;  // [E*E + 1]
    int32_t* pos;         // [N]   token -> row in pair order
    int32_t* tok;         // [N]   row in pair order -> token
    int32_t* rowmap;      // [N*K] expert-major dispatched row -> pair-order row | (route << 30)
    I64bGroup* groups;    // [E*E] one GEMM group per pair (w_row_base = e0*H, pad0 = e1*H)
    int32_t* ticket;      // self-resetting last-CTA counter
    size_t bytes;
};
I64bTriWs i64b_tri_ws_carve(void* base, int64_t n, int top_k, int num_experts);
// after i64b_run_plan (reads the plan's packed expert ids): 1 launch
int i64b_run_tri_plan(const I64bPlanWs& plan, const I64bTriWs& tri, int64_t n, int num_experts, int hidden,
                      cudaStream_t stream);

struct I64bDispatchArgs {
    const void* x;              // [N, H] bf16 (fused path) or raw rows
    int64_t n;
    int hidden;
    int top_k, num_experts;
    int quantize;               // 1: per-row int8 quantisation fused into the scatter; 2: per-row FP8 E4M3 (scale = absmax / 448)
    const float* preq_scale;    // non-null: x is ALREADY int8 [N, H] with these per-token scales (x_preq from the fused
                                // RMSNorm+quant): rows and scales are copied, nothing is re-quantised (quantize must be 0)
    int shared_copy;            // 1: also write row N*K + t (token order) for the shared expert
    void* xs;                   // [N*K (+N), H] bf16 or int8
    float* xs_scale;            // [N*K (+N)] (quantize)
    // optional reference-convention outputs
    int32_t* routes_out;        // [K, N]
    int32_t* scatter_idx_out;   // [K, N]
    int32_t* dest_row_out;      // [K, N]
    int32_t* src_tok_out;       // [N*K] optional: token of every dispatched row (GEMM1 gathers its A rows through it)
    int tri;                    // 1: also fill tw.pos / tw.tok / tw.rowmap (token-major layout, top_k == 2)
    int phase;                  // 0: indices and rows in one launch; 1: the index outputs only (routes / dest rows / token-major
                                // maps); 2: the row copies only — the copy then runs on a second stream under the shared expert's
                                // GEMM1, which needs the indices but not the dispatched rows
    I64bTriWs tw;
};
int i64b_run_dispatch(const I64bDispatchArgs& a, const I64bPlanWs& ws, cudaStream_t stream);

// q14 int32 [rows, cols] (+ per-(row, tile) |max| partials) -> int8 rows + fp32 scale (bit-exact re-quantisation
// of silu_multiply_integer's fp32 output, core/quantization.py:103-116)
// (two row sets per launch: routed assignments and the shared expert's rows; either may be empty)
int i64b_run_requant_q14(const int32_t* q14, const int32_t* amax_part, int amax_ld, int n_tiles, int8_t* out,
                         float* scale, int64_t rows, int cols, const int32_t* q14_b, const int32_t* amax_part_b,
                         int amax_ld_b, int n_tiles_b, int8_t* out_b, float* scale_b, int64_t rows_b, int cols_b,
                         cudaStream_t stream, const I64bGroup* groups = nullptr, int num_groups = 0,
                         const struct I64bTriWs* tri = nullptr);
// `tri` (W8A8 fused down projection): the int8 rows go to the token-major [rows_b, 2*cols + cols_b] buffer `out` (routed
// row r -> row tri->rowmap[r], K segment = its bit 30; shared row t -> row tri->pos[t], last segment) and the scales to
// `scale` as [3, rows_b] by pair-order row; out_b / scale_b are ignored
// `groups` (device): the first row set is the rows of these GEMM groups instead of rows [0, rows) — `rows` is then only
// an upper bound used to size the grid

// bf16 SwiGLU rows [rows, cols] -> FP8 E4M3 rows + one fp32 scale per row (quantize_activations_fp8, core/fp8.py:113-129);
// two row sets per launch (routed assignments, shared expert), either may be empty
int i64b_run_requant_fp8(const void* h, uint8_t* out, float* scale, int64_t rows, int cols, const void* h_b, uint8_t* out_b,
                         float* scale_b, int64_t rows_b, int cols_b, cudaStream_t stream,
                         const struct I64bTriWs* tri = nullptr);      // tri: token-major scatter, as i64b_run_requant_q14

struct I64bCombineArgs {
    const float* y;             // [N*K + N, H] fp32 per-assignment outputs (shared rows last)
    const float* y_shared;      // optional separate [N, H] shared-expert outputs (else rows N*K.. of y)
    const int32_t* dest_row;    // [K, N]
    int64_t n;
    int hidden, top_k, has_shared, has_routed;
    float w[I64B_MAX_TOPK];
    float routed_scale, shared_scale;
    void* out;
    int out_dtype;
};
int i64b_run_combine(const I64bCombineArgs& a, cudaStream_t stream);
