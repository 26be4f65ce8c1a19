/* i64b.h — C ABI of the B200-native token-routed MoE FFN block for vllm-i64.
 *
 * Drop-in boundary: every entry point takes plain device pointers, sizes and an
 * explicit CUDA stream; returns 0 or a negative I64B_ERR_* code (text via
 * i64b_last_error(), thread-local); never throws, never allocates device memory,
 * never synchronises, never touches the legacy default stream unless the caller
 * passes stream 0.  All outputs and workspaces are caller-allocated (query sizes
 * with the *_bytes functions).  Each function cites the reference interface it
 * replaces (paths relative to the Complexity-ML/vllm-i64 checkout).
 *
 * The reference's own native conventions this improves on: csrc/i64_router.cu:200-268
 * and csrc/i64_expert_dispatch.cu:144-233 (extern "C" entry points that throw,
 * cudaMallocAsync temporaries and a D2H sync per call).
 */
#ifndef I64B_H_
#define I64B_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* i64b_stream_t; /* cudaStream_t */

enum {
    I64B_OK = 0,
    I64B_ERR_ARG = -1,       /* bad argument / unsupported shape */
    I64B_ERR_CUDA = -2,      /* CUDA runtime or driver error */
    I64B_ERR_WORKSPACE = -3, /* workspace too small */
    I64B_ERR_STATE = -4      /* layer not ready (weights / table not loaded) */
};

enum { I64B_BF16 = 0, I64B_F32 = 1, I64B_F16 = 2 };
enum { I64B_MODE_BF16 = 0, I64B_MODE_W8A8 = 1, I64B_MODE_FP8 = 2 };

const char* i64b_last_error(void);
int i64b_version(void);
/* number of kernels launched by this thread through the library since the last reset */
int64_t i64b_launch_count(void);
void i64b_launch_count_reset(void);
/* host pointer to the 2049-entry Q7 SiLU table (round(silu(i/128)*128), i in [-1024,1024]) used by the W8A8
 * SwiGLU epilogue; identical to _SILU_LUT of layers/integer_activations.py:28-36. */
const int16_t* i64b_silu_lut_q7(void);

/* ---------------------------------------------------------------------------
 * Routing  (a1, a6)
 * ------------------------------------------------------------------------- */

/* expert = token_id & (E-1), int32.  Replaces route_tokens (csrc/i64_ops_binding.cu:32-65),
 * i64_route_tokens_kernel (csrc/i64_router.cu:51-70), kernels/i64_ops.py:17-34. */
int i64b_route_modulo(const int64_t* token_ids, int32_t* expert_ids, int64_t n, int num_experts,
                      i64b_stream_t stream);

/* Pack the checkpoint's topk_token_to_expert [K,V] int64 table to uint8 (load time).
 * Source of the table: core/loader.py:374-407. */
int i64b_pack_table(const int64_t* table_i64, uint8_t* table_u8, int top_k, int vocab, int num_experts,
                    i64b_stream_t stream);

/* routes[k,t] = table[k, clamp(token_ids[t],0,V-1)]; token_ids == NULL -> (0+k)%E.
 * Replaces TokenRoutedMLP.route (layers/token_routed_mlp.py:113-130).  routes is int64 [K,N]. */
int i64b_route_table(const int64_t* token_ids, const uint8_t* table_u8, int64_t* routes, int64_t n, int vocab,
                     int top_k, int num_experts, i64b_stream_t stream);

/* ---------------------------------------------------------------------------
 * Dispatch plan / scatter / gather  (a7, a8)
 * ------------------------------------------------------------------------- */

size_t i64b_plan_workspace_bytes(int64_t n, int top_k, int num_experts);

/* Deterministic (stable-by-token-index) dispatch plan for K routes at once.
 *   expert_ids      int32 [K,N] (device)
 *   scatter_indices int32 [K,N] per-route destination row (== stable argsort inverse), nullable
 *   counts/offsets  int32 [K,E] per-route bincount and exclusive scan, nullable
 *   dest_row        int32 [K,N] row in the merged expert-major layout, nullable
 *   group_rows      int32 [E]   rows per expert in the merged layout, nullable
 * Replaces scatter_by_expert's argsort/bincount/cumsum (csrc/i64_ops_binding.cu:179-204),
 * i64_prefix_sum_kernel + atomic slot claiming (csrc/i64_router.cu:94-134,171-184). */
int i64b_dispatch_plan(const int32_t* expert_ids, int top_k, int64_t n, int num_experts,
                       int32_t* scatter_indices, int32_t* counts, int32_t* offsets, int32_t* dest_row,
                       int32_t* group_rows, void* workspace, size_t workspace_bytes, i64b_stream_t stream);

/* The same plan, same kernels, same five outputs (each nullable) as i64b_dispatch_plan, for callers that cannot supply a
 * workspace.  Nothing in the reference corresponds to this entry point; it exists for two kinds of host:
 *   - a serving process that gave every free byte of HBM to its KV-cache pool at start-up, where a cudaMalloc issued
 *     while serving fails (or stalls all streams of the device while it is mapped);
 *   - an FFI host that only receives device pointers from elsewhere and binds no device allocator.
 * The scratch is an array compiled into the library (device global memory, I64B_DISPATCH_PLAN_STATIC_BYTES per GPU that
 * loads libi64b.so, held until the process exits).  The call performs no cudaMalloc / cudaFree and uses no allocating
 * library routine; inputs and outputs remain the caller's.
 *   Limit: i64b_plan_workspace_bytes(n, top_k, num_experts) <= I64B_DISPATCH_PLAN_STATIC_BYTES (about 250 000
 *     (token, route) assignments); anything larger returns I64B_ERR_WORKSPACE with the two sizes in i64b_last_error().
 *   First call on each device: looks the array's device address up (cudaGetSymbolAddress, with the caller's current
 *     device) and remembers it.  Make that first call during set-up, with the right device current and before any
 *     stream capture begins on that device; later calls only launch kernels.
 *   Sharing: there is ONE array per device and every call uses it from offset 0.  Calls on the same stream are ordered
 *     and safe.  Calls that can overlap on the GPU (different streams without an event between them, or different
 *     host threads on different streams) race on the scratch and produce garbage plans — serialise them, or use
 *     i64b_dispatch_plan with one workspace per stream.  The array keeps whatever the last call wrote; the plan never
 *     reads a word before writing it, so nothing is (or needs to be) cleared between calls. */
#define I64B_DISPATCH_PLAN_STATIC_BYTES (1 << 20)
int i64b_dispatch_plan_static(const int32_t* expert_ids, int top_k, int64_t n, int num_experts,
                              int32_t* scatter_indices, int32_t* counts, int32_t* offsets, int32_t* dest_row,
                              int32_t* group_rows, i64b_stream_t stream);

/* out[dest[t], :] = x[t, :] for row_bytes-wide rows (row_bytes % 16 == 0).
 * Replaces i64_scatter_kernel (csrc/i64_router.cu:94-134), atomic_scatter / fused_route_scatter row copies. */
int i64b_scatter_rows(const void* x, const int32_t* dest, void* out, int64_t n, int64_t row_bytes,
                      i64b_stream_t stream);

/* out[t, :] = src[idx[t], :].  Replaces i64_gather_kernel (csrc/i64_router.cu:150-165),
 * gather_by_expert (csrc/i64_ops_binding.cu:207-209). */
int i64b_gather_rows(const void* src, const int32_t* idx, void* out, int64_t n, int64_t row_bytes,
                     i64b_stream_t stream);

/* Weighted un-permute combine: out[t,:] = routed_scale * sum_k w[k] * y[dest_row[k,t],:] + shared_scale *
 * y_shared[t,:] in the reference's fp32 operation order (layers/token_routed_mlp.py:156-185); y fp32 [rows,H],
 * y_shared nullable.  Used stand-alone by the expert-parallel path (rows come back from the return all-to-all). */
int i64b_combine(const float* y, const int32_t* dest_row, const float* y_shared, int top_k, const float* route_weight,
                 float routed_scale, float shared_scale, void* out, int out_dtype, int64_t n, int hidden,
                 i64b_stream_t stream);

/* ---------------------------------------------------------------------------
 * Elementwise / quantisation ops  (a9, a12)
 * ------------------------------------------------------------------------- */

/* out[n,i] = silu(gate_up[n,i]) * gate_up[n,inter+i]; dtype in {BF16,F32,F16}.
 * Replaces silu_hadamard (csrc/i64_ops_binding.cu:74-172), i64_silu_hadamard_kernel
 * (csrc/i64_expert_dispatch.cu:100-122), kernels/triton/I64_fused_silu_mul.py. */
int i64b_silu_hadamard(const void* gate_up, void* out, int64_t n, int inter, int dtype, i64b_stream_t stream);

/* Per-row symmetric INT8: s = max(absmax,1e-8)/127, q = clamp(rint(x/s)).  dtype of x in {BF16,F32,F16};
 * arithmetic is IEEE fp32 (bit-exact with core/quantization.py:80-116).
 * Replaces I64_quantize_int8 / I64_quantize_perchannel_int8 (csrc/I64_quantize.cu:41-92,118-166,173-243). */
int i64b_quantize_rows_int8(const void* x, int dtype, int8_t* q, float* scale, int64_t rows, int64_t cols,
                            i64b_stream_t stream);
/* out_f32[r,c] = q[r,c] * scale[r].  Replaces I64_dequantize_int8 (csrc/I64_quantize.cu:99-111). */
int i64b_dequantize_int8(const int8_t* q, const float* scale, float* out, int64_t rows, int64_t cols,
                         i64b_stream_t stream);

/* Fused residual-add + RMSNorm (+ per-token INT8 quantisation) — the step in front of the block:
 *     h = residual + x ; y = rmsnorm(h) * weight ; (q, scale) = quantize_rows_int8(y)
 * in ONE pass over the rows.  x, residual (nullable), weight [H] have dtype BF16 or F32; hidden_out (nullable, needs
 * residual) receives h — the next residual; normed_out (nullable) y in the same dtype; q_out/scale_out (nullable) the
 * int8 form the W8A8 block consumes as x_preq (i64b_moe_forward_preq).  bf16 rounds where the reference's eager path
 * rounds (layers/rmsnorm.py:86-90: "(x.float()*norm).type_as(x) * weight"); F32 is the contract of the reference's own
 * kernel: y = x*rrms*weight in fp32.  Replaces models/complexity_deep/model.py:358-361 (residual add + post-attention
 * norm as separate elementwise kernels), I64_rmsnorm_forward / I64_rmsnorm_quant_forward (csrc/I64_rmsnorm.cu:41-190),
 * kernels/triton/I64_fused_rmsnorm_quant.py. */
int i64b_add_rmsnorm(const void* x, const void* residual, const void* weight, int dtype, float eps, void* hidden_out,
                     void* normed_out, int8_t* q_out, float* scale_out, int64_t n, int hidden, i64b_stream_t stream);

/* ---------------------------------------------------------------------------
 * The block as one object  (a2, a3, a4, a5, a12)
 * ------------------------------------------------------------------------- */

typedef struct {
    int32_t hidden_size;        /* H, multiple of 64 (128 in W8A8 mode) */
    int32_t num_experts;        /* E in [1,64] (0 = dense FFN: only the "shared" SwiGLU runs) */
    int32_t expert_inter;       /* I_e per expert, multiple of 16 */
    int32_t shared_inter;       /* S, 0 = no shared expert */
    int32_t vocab_size;         /* V */
    int32_t top_k;              /* K in [1,4] */
    int32_t mode;               /* I64B_MODE_BF16 | I64B_MODE_W8A8 | I64B_MODE_FP8 */
    int32_t reserved;
    float route_weight[4];      /* w0 = primary, others (1-w0)/(K-1)  (layers/token_routed_mlp.py:53-58,158) */
    float routed_scale;         /* routed_output_scale or routed_output_gate (:174-185) */
    float shared_scale;         /* shared_output_scale or shared_output_gate */
} i64b_moe_config;

typedef struct i64b_layer i64b_layer;

int i64b_moe_create(const i64b_moe_config* cfg, i64b_layer** out);
void i64b_moe_destroy(i64b_layer* layer);
int i64b_moe_set_scales(i64b_layer* layer, const float* route_weight, int top_k, float routed_scale,
                        float shared_scale);

/* bytes of device memory the caller must provide for the packed (tcgen05-friendly, K-major) weights */
size_t i64b_moe_packed_bytes(const i64b_layer* layer);

/* Repack reference-layout bf16 weights (layers/token_routed_mlp.py:66-82): gate,up [E,H,I_e];
 * down [E,I_e,H]; shared_gate,shared_up [S,H]; shared_down [H,S] (nn.Linear (out,in)), into `packed`.
 * Replaces load_full_weights / fuse_shared_gate_up / materialize_expert_linears (:189-258).
 * Shared pointers may be NULL when shared_inter == 0; expert pointers NULL when num_experts == 0. */
int i64b_moe_load_bf16(i64b_layer* layer, const void* gate, const void* up, const void* down,
                       const void* shared_gate, const void* shared_up, const void* shared_down, void* packed,
                       i64b_stream_t stream);

/* W8A8: weights already quantised per out-row (quantize_int8, core/quantization.py:80-91) in the
 * materialised (out,in) form: w13 [E,2I_e,H] int8 (+[E,2I_e] fp32 scales), w2 [E,H,I_e] (+[E,H]);
 * shared sh13 [2S,H], sh2 [H,S].  Copied into `packed`. */
int i64b_moe_load_w8a8(i64b_layer* layer, const int8_t* w13_q, const float* w13_s, const int8_t* w2_q,
                       const float* w2_s, const int8_t* sh13_q, const float* sh13_s, const int8_t* sh2_q,
                       const float* sh2_s, void* packed, i64b_stream_t stream);

/* FP8 (f4): the block with E4M3 weights and dynamically quantised E4M3 activations, fp32 accumulation on the tensor cores
 * (tcgen05 kind::f8f6f4).  Semantics of core/fp8.py: weights quantised per out-row (quantize_fp8, :75-110: scale =
 * absmax / 448), activations per token row (quantize_activations_fp8, :113-129), every projection
 * y = (x_q . w_q^T) * (x_scale x w_scale) (fp8_linear / fp8_fused_gate_up via torch._scaled_mm, :137-245), the SwiGLU rows
 * rounded to bf16 and re-quantised per row in front of the down projection (layers/dense_mlp.py:75-83).  Same layouts as
 * i64b_moe_load_w8a8 with E4M3 bytes (torch.float8_e4m3fn viewed as uint8) in place of the int8 ones. */
int i64b_moe_load_fp8(i64b_layer* layer, const uint8_t* w13_q, const float* w13_s, const uint8_t* w2_q,
                      const float* w2_s, const uint8_t* sh13_q, const float* sh13_s, const uint8_t* sh2_q,
                      const float* sh2_s, void* packed, i64b_stream_t stream);

/* uint8 routing table [K,V] (device, from i64b_pack_table); kept by pointer. */
int i64b_moe_set_table(i64b_layer* layer, const uint8_t* table_u8);

size_t i64b_moe_workspace_bytes(const i64b_layer* layer, int64_t max_tokens);

/* out[N,H] = routed_scale * sum_k w_k * expert_{route_k(t)}(x[t]) + shared_scale * shared(x[t]).
 * x: bf16 [N,H] device; token_ids: int64 [N] device or NULL; out: out_dtype in {BF16,F32}.
 * Replaces TokenRoutedMLP.forward (layers/token_routed_mlp.py:154-187) and, with num_experts == 0,
 * DenseSwiGLUMLP.forward / DenseMLP.forward (models/complexity_deep/model.py:250-255, layers/dense_mlp.py:46-97). */
int i64b_moe_forward(i64b_layer* layer, const void* x, const int64_t* token_ids, void* out, int out_dtype,
                     int64_t n, void* workspace, size_t workspace_bytes, i64b_stream_t stream);
/* The same forward on activations the caller has ALREADY quantised (x_preq of the reference: layers/dense_mlp.py:46-58,
 * core/quantization.py:124,157-158 — produced by the fused RMSNorm+quant, i64b_add_rmsnorm): x_q int8 [N,H], x_scale fp32
 * [N].  W8A8 layers only.  Skips the absmax/quantise pass of the dispatch (rows and scales are copied) and the shared
 * expert reads x_q in place; the result is bit-identical to i64b_moe_forward on the rows x_q was quantised from. */
int i64b_moe_forward_preq(i64b_layer* layer, const int8_t* x_q, const float* x_scale, const int64_t* token_ids, void* out,
                          int out_dtype, int64_t n, void* workspace, size_t workspace_bytes, i64b_stream_t stream);
/* Contract of the single-launch decode kernel (bf16 layers, n <= 1024 — i64b_moe_uses_fused): its CTAs wait on one
 * another (GEMM1 -> GEMM2 -> combine inside one launch) and it is launched with one CTA per SM, so it completes only if
 * the whole grid becomes resident.  ONE such launch may be in flight per device at a time: issue decode-sized forwards of
 * one device on ONE stream (as the reference's engine does: one step at a time, cpu/engine.py:195-198), or set
 * I64B_DEBUG_NO_FUSED on layers that run on a second stream / next to another process under MPS (they then take the
 * kernel chain, which has no inter-CTA waits).  Every in-kernel wait is bounded (__trap after ~2 s), never a hang.
 * The layer handle is not internally locked: one forward per layer at a time. */

/* 1 if i64b_moe_forward on n tokens takes the single-launch fused decode kernel (bf16, n <= 1024; env
 * I64B_FUSED=0 disables it), else 0 (plan / dispatch / GEMM1 / GEMM2 / combine as separate launches). */
int i64b_moe_uses_fused(const i64b_layer* layer, int64_t n);
/* The same question for a shape alone (pure host arithmetic, no device needed): bf16 block of `n` tokens with these
 * widths.  0 when the single-launch kernel's limits are exceeded (n <= 1024, widths <= 4096, hidden % 64 == 0, at most
 * 8 work units per SM in its static work list, ...): such shapes take the kernel chain. */
int i64b_fused_decode_supported(int n, int hidden, int expert_inter, int shared_inter, int num_experts, int top_k);
/* 1 if the kernel chain on n tokens finishes the block inside the down-projection's epilogue (no separate combine
 * launch), else 0.  Meaningless (0) when i64b_moe_uses_fused is 1. */
int i64b_moe_fuses_combine(const i64b_layer* layer, int64_t n);
/* Which form of the token-major fused down projection a plain forward of n tokens takes: 0 none (GEMM2 + combine),
 * 1 three accumulators folded in the epilogue (bit-identical to GEMM2 + combine_kernel), 2 one accumulator over
 * K = 2*I_e + S with the scales — all powers of two — folded into the SwiGLU rows. */
int i64b_moe_tri_form(const i64b_layer* layer, int64_t n);

/* Run only the selected stages of the forward on the workspace contents left by an earlier full forward with the
 * same (n, workspace): used by bench.py to time each kernel in isolation (roofline per stage). */
enum {
    I64B_STAGE_PLAN = 1, I64B_STAGE_DISPATCH = 2, I64B_STAGE_GEMM1 = 4, I64B_STAGE_REQUANT = 8,
    I64B_STAGE_GEMM2 = 16, I64B_STAGE_COMBINE = 32, I64B_STAGE_ALL = 0xFF
};
int i64b_moe_forward_stages(i64b_layer* layer, const void* x, const int64_t* token_ids, void* out, int out_dtype,
                            int64_t n, void* workspace, size_t workspace_bytes, unsigned stage_mask,
                            i64b_stream_t stream);

/* Same block with the routes supplied by the caller: routes int32 [K,N] expert ids (device).  This is the
 * form the reference's kernels package uses — fused_token_routed_forward(x, gate_up, down, expert_ids, ...)
 * (kernels/fused_experts.py:29-150), i64_expert_forward / i64_full_pipeline (kernels/i64_ops.py:100-193),
 * i64_expert_dispatch (csrc/i64_expert_dispatch.cu:144-233). */
int i64b_moe_forward_routed(i64b_layer* layer, const void* x, const int32_t* routes, int top_k,
                            const float* route_weight, float routed_scale, float shared_scale, int use_shared,
                            void* out, int out_dtype, int64_t n, void* workspace, size_t workspace_bytes,
                            i64b_stream_t stream);
size_t i64b_moe_workspace_bytes_routed(const i64b_layer* layer, int64_t max_tokens, int top_k);

/* stream-ordered device-to-device copy (lets a ctypes host read the debug taps without a torch extension) */
int i64b_copy_device(void* dst, const void* src, size_t bytes, i64b_stream_t stream);

/* Same call with HOST buffers (pinned or pageable): H2D of x/token_ids, forward, D2H of out, all
 * enqueued on `stream`; x_dev/ids_dev/out_dev are caller-provided device staging buffers. */
int i64b_moe_forward_host(i64b_layer* layer, const void* x_host, const int64_t* token_ids_host, void* out_host,
                          int out_dtype, int64_t n, void* x_dev, int64_t* ids_dev, void* out_dev,
                          void* workspace, size_t workspace_bytes, i64b_stream_t stream);

/* Debug taps into the workspace of the LAST forward on this layer (device pointers, valid until the
 * next forward): lets the parity tests compare dispatch plan, INT8 activations and INT32 accumulators
 * bit-for-bit.  `what` is one of the I64B_TAP_* ids; returns NULL if not applicable. */
enum {
    I64B_TAP_ROUTES = 0,      /* int32 [K,N] expert ids */
    I64B_TAP_DEST_ROW = 1,    /* int32 [K,N] merged-layout row */
    I64B_TAP_SCATTER_IDX = 2, /* int32 [K,N] per-route row */
    I64B_TAP_COUNTS = 3,      /* int32 [K,E] */
    I64B_TAP_OFFSETS = 4,     /* int32 [K,E] */
    I64B_TAP_XS = 5,          /* dispatched activations [N*K (+ N in W8A8: quantised token-order copy), H] (bf16 or int8) */
    I64B_TAP_XS_SCALE = 6,    /* fp32 [N*K + N] (W8A8) */
    I64B_TAP_INTER_R = 7,     /* routed intermediate [N*K, I_e] (bf16 or int8) */
    I64B_TAP_INTER_S = 8,     /* shared intermediate [N, S] */
    I64B_TAP_INTER_SCALE_R = 9,
    I64B_TAP_INTER_SCALE_S = 10,
    I64B_TAP_Q14_R = 11,      /* int32 [N*K, I_e] (W8A8) */
    I64B_TAP_Q14_S = 12,      /* int32 [N, S] */
    I64B_TAP_Y = 13,          /* fp32 [N*K + N, H] per-assignment expert outputs */
    I64B_TAP_GROUP_ROWS = 14  /* int32 [E] */
};
void* i64b_moe_tap(i64b_layer* layer, int what);

/* Debug switches of a layer.  I64B_DEBUG_KEEP_TAPS: the down-projection kernel finishes the block inside the shared
 * expert's tiles (fused combine) and then has no reason to store the shared expert's own fp32 rows; with this flag it
 * stores them as well, so I64B_TAP_Y is complete (rows N*K.. = shared expert).  The block output is identical. */
enum {
    I64B_DEBUG_KEEP_TAPS = 1,
    I64B_DEBUG_FUSE_COMBINE = 2,  /* force the fused combine (default: dense FFN only) */
    I64B_DEBUG_NO_FUSED = 4,      /* never take the single-launch decode kernel on this layer (it must own the GPU: its
                                     CTAs wait on one another, so it may not share the device with another launch of
                                     itself — callers that run this layer on a second stream set this) */
    I64B_DEBUG_TRI = 8,           /* take the token-major fused down projection whenever the shape allows (top-2 bf16
                                     blocks, E*E <= 64; default: only when the expert-pair groups are tall enough) */
    I64B_DEBUG_NO_TRI = 16,       /* never take it (per-assignment fp32 rows + combine_kernel) */
    I64B_DEBUG_NO_OVERLAP = 64,   /* with env I64B_OVERLAP=1 (off by default: measured slower) the dispatch row copies of bf16
                                     blocks with a shared expert run, from 4096 tokens on, on the layer's second stream under
                                     the shared expert's GEMM1; this flag keeps them on the caller's stream.  Results are
                                     identical either way */
    I64B_DEBUG_GATHER = 128,      /* bf16 blocks big enough for the CTA-pair GEMM1: skip the dispatch row copies, GEMM1 gathers
                                     the routed A rows from x (tile::gather4 through the dispatch indices).  Identical results,
                                     measured much slower (the gather starves GEMM1's operand ingest): off unless this flag
                                     or env I64B_GATHER=1 */
    I64B_DEBUG_TRI_FOLD = 32      /* fused down projection: keep the three-accumulator form (bit-identical to GEMM2 +
                                     combine_kernel) even when the scales are powers of two and the single-accumulator
                                     form (scales folded into the SwiGLU rows, one sum over K = 2*I_e + S) would be taken */
};
/* auto selection of the fused down projection: at least this many tokens per unordered expert pair, E(E-1)/2 pairs.
 * Batch invariance: the choice depends on n, and the single-accumulator form sums a token's three contributions in one
 * fp32 accumulation where every other form rounds three partial sums — so with power-of-two scales the low bf16 bit of a
 * token's output can differ between a batch below and a batch above I64B_TRI_MIN_ROWS_ONEACC rows per pair (<= 1 bf16 ulp,
 * <= 2e-5 of max|out| in fp32).  A caller that needs the same bits for a token whatever batch it arrives in sets
 * I64B_DEBUG_TRI_FOLD (or env I64B_TRI_ONEACC=0): all remaining forms are bit-identical to one another. */
#define I64B_TRI_MIN_ROWS_ONEACC 288
#define I64B_TRI_MIN_ROWS_FOLD 512
int i64b_moe_set_debug(i64b_layer* layer, int flags);

/* Cross-layer weight prefetch (decode).  `next` is the layer the engine runs after `layer` (the decoder stack calls
 * its MLPs in a fixed order: models/complexity_deep/model.py:478-486); the fused decode kernel of `layer` then pulls
 * next's packed weights into L2 while it waits on its own dependencies.  Results are unaffected.  next == NULL turns
 * it off; `next` must outlive `layer` or be cleared first.  OFF unless env I64B_PREFETCH is 1 (prefetch at kernel
 * entry) or 2 (after the layer's own weight stream): measured on B200 it does not pay (24.7 us/layer without,
 * 26.2 / 25.3 us with — the layer is bound by its dependency chain, not by where the weights come from). */
int i64b_moe_set_next(i64b_layer* layer, const i64b_layer* next);

/* GEMM launches of this layer use at most max_ctas CTAs (0 = one per SM, the default).  The expert-parallel host leaves
 * a few SMs to the NVLink exchange kernels that run next to the shared expert's GEMMs on another stream: a 210 KB GEMM
 * CTA per SM leaves no shared memory for them otherwise, and they would queue behind the whole GEMM.  Results do not
 * depend on it (the tile -> CTA assignment never enters the arithmetic). */
int i64b_moe_set_cta_limit(i64b_layer* layer, int max_ctas);

/* Bring-up only (env I64B_DEBUG_TIMELINE=1): after replaying a CUDA graph of fused decode layers, print the per-CTA
 * globaltimer stamps of the last two layers to stderr.  Synchronises the device.  Returns 0 when no timeline exists. */
int i64b_debug_timeline_dump(void);

/* ---------------------------------------------------------------------------
 * Expert parallelism over NVLink peer memory  (a14)
 * Replaces TokenRoutedMLP._forward_ep (layers/token_routed_i64/layer.py:209-298): counts all-to-all + .tolist() +
 * argsort + 2 x all_to_all_single each way.  One process per GPU; every rank allocates its buffers with
 * i64b_ep_alloc, exports them with i64b_ipc_export, exchanges the 64-byte handles through its own control plane
 * (torch.distributed.all_gather_object in the Python host, once) and maps the peers' with i64b_ipc_import.
 *
 * Layout: every receive buffer is cut into `size` regions of cap_src rows, one per SOURCE rank.  A rank plans only
 * its OWN tokens (routing is a pure function of layer and token id); in its expert-major plan the experts of owner d
 * are one contiguous slab, which it writes, row by row, into its region of d's buffer, and it publishes the slab's
 * per-expert row counts in d's flags page.  No rank ever needs another rank's token ids: no id all-gather, no global
 * plan, no split-size exchange, and ranks may bring different numbers of tokens.
 * Per forward (all on one stream, no host synchronisation, no collective library call; CUDA-graph capturable):
 *   i64b_ep_dispatch      local plan + rows (bf16, or int8 + fp32 scale in W8A8 mode) into the owners' buffers + counts
 *   i64b_ep_barrier       "my rows have left" / every peer's rows have arrived   (the home rank's shared expert runs on
 *                         a second stream meanwhile)
 *   i64b_moe_forward_dispatched   group tables from the counts, then this rank's experts on its receive buffer ->
 *                         per-assignment rows in this rank's y buffer (same row numbering as the receive buffer)
 *   i64b_ep_barrier
 *   i64b_ep_combine       the home rank's weighted combine, reading the K rows of every token straight from the
 *                         owners' y buffers over NVLink (the return exchange is fused into the combine)
 * ------------------------------------------------------------------------- */
#define I64B_EP_MAX_RANKS 8
#define I64B_EP_FLAGS_INTS 1024    /* ints per flags page: [0,8) barrier epochs by source rank; [64, 64 + 8*64) row counts
                                      [source rank][local expert]; zero-initialised by i64b_ep_alloc */
typedef struct i64b_ep_ctx {
    int32_t rank, size;            /* EP rank / group size */
    int32_t n_local;               /* tokens THIS rank brings to this call (ranks may differ); n_local * top_k <= cap_src */
    int32_t hidden, top_k, num_experts, vocab;
    int32_t y_dtype;               /* dtype the expert outputs travel back in: I64B_BF16 (the reference's EP path returns
                                      expert outputs in the model dtype) or I64B_F32 (bit-identical to the single-GPU block) */
    int32_t x_mode;                /* I64B_MODE_BF16: rows travel as bf16; I64B_MODE_W8A8: quantised once on the home rank,
                                      int8 rows + one fp32 scale per row travel (half the dispatch bytes) */
    int32_t comm_ctas;             /* > 0: the dispatch / combine kernels use this many CTAs (one per SM, large shared memory),
                                      leaving the other SMs to GEMMs that run next to them on other streams (the layers
                                      doing so carry i64b_moe_set_cta_limit(layer, SMs - comm_ctas)); 0: every SM */
    int64_t cap_src;               /* rows of the region each source rank owns in every receive buffer */
    void* peer_xs[I64B_EP_MAX_RANKS];          /* [size * cap_src, H] bf16 | int8 receive buffer of rank d, mapped here */
    float* peer_xs_scale[I64B_EP_MAX_RANKS];   /* [size * cap_src] fp32 row scales (W8A8 transport), else unused */
    void* peer_y[I64B_EP_MAX_RANKS];           /* [size * cap_src, H] y_dtype per-assignment outputs of rank d's experts */
    int32_t* peer_flags[I64B_EP_MAX_RANKS];    /* [I64B_EP_FLAGS_INTS] flags page of rank d */
    int32_t* state;                /* local, zero-initialised: [0] barrier epoch */
} i64b_ep_ctx;

int i64b_ep_alloc(size_t bytes, void** out);               /* cudaMalloc + zero fill (setup only) */
int i64b_ep_free(void* p);
int i64b_ipc_export(void* p, unsigned char* handle64);     /* cudaIpcGetMemHandle */
int i64b_ipc_import(const unsigned char* handle64, void** out);
int i64b_ipc_close(void* p);
/* sized for cap_src / top_k tokens, whatever n_local is */
size_t i64b_ep_workspace_bytes(const i64b_ep_ctx* ctx);
/* The plan half of i64b_ep_dispatch alone.  Routing is a pure function of (layer, token id), so an engine that knows the
 * step's token ids can plan EVERY layer of the step up front on a side stream (the scheduler pre-route of
 * engine/i64_scheduler.py:293-298, here per layer because every layer has its own table): i64b_ep_dispatch called with
 * token_ids == NULL then uses the plan found in the workspace (same ctx->n_local) and only moves rows. */
int i64b_ep_plan(const i64b_ep_ctx* ctx, const int64_t* token_ids, const uint8_t* table_u8, void* workspace,
                 size_t workspace_bytes, i64b_stream_t stream);
/* x: [n_local, H] bf16 (this rank's tokens); token_ids: int64 [n_local] (this rank's), or NULL after i64b_ep_plan;
 * table_u8: [K, V] */
int i64b_ep_dispatch(const i64b_ep_ctx* ctx, const void* x, const int64_t* token_ids, const uint8_t* table_u8,
                     void* workspace, size_t workspace_bytes, i64b_stream_t stream);
/* out[n_local, H] = routed_scale * sum_k w_k * y[owner(k,t)][row(k,t)] + shared_scale * y_shared[t] (y_shared: fp32
 * [n_local, H] or NULL); same operation order as i64b_combine */
int i64b_ep_combine(const i64b_ep_ctx* ctx, void* ep_workspace, const float* y_shared, const float* route_weight,
                    float routed_scale, float shared_scale, void* out, int out_dtype, i64b_stream_t stream);
/* flag barrier in peer memory: signal (system-scope release after this rank's earlier peer stores) then wait for every
 * peer's signal of the same epoch; device-resident epochs, so a captured graph replays correctly */
int i64b_ep_barrier(const i64b_ep_ctx* ctx, i64b_stream_t stream);
int i64b_ep_signal(const i64b_ep_ctx* ctx, i64b_stream_t stream);    /* the two halves, for callers that put work between */
int i64b_ep_wait(const i64b_ep_ctx* ctx, i64b_stream_t stream);
/* `layer` holds THIS rank's experts (num_experts / size of them, no shared expert; bf16 layer for bf16 transport, W8A8
 * layer for int8 transport); ep_workspace is the buffer i64b_ep_dispatch used.  Runs gate|up + SwiGLU (+ re-quant) +
 * down on the receive buffer into ctx->peer_y[rank]. */
size_t i64b_moe_dispatched_workspace_bytes(const i64b_layer* layer, const i64b_ep_ctx* ctx);
int i64b_moe_forward_dispatched(i64b_layer* layer, const i64b_ep_ctx* ctx, void* ep_workspace, void* workspace,
                                size_t workspace_bytes, i64b_stream_t stream);

/* ---------------------------------------------------------------------------
 * Stand-alone W8A8 / weight-only GEMM ops  (a12, a13)
 * ------------------------------------------------------------------------- */

/* FP8 (f4).  Per-row E4M3 quantisation of [rows, cols] bf16 / fp16 / fp32 rows: scale = max(absmax, 1e-12) / 448,
 * q = e4m3((x / scale).clamp(+-448)) in IEEE fp32, round to nearest even — quantize_activations_fp8 and
 * quantize_fp8(per_channel=True), core/fp8.py:75-129; bit-exact with them. */
int i64b_quantize_rows_fp8(const void* x, int dtype, uint8_t* q, float* scale, int64_t rows, int64_t cols,
                           i64b_stream_t stream);
/* out[m,n] = (x_q[m,:] . w_q[n,:]) * (x_scale[m] * w_scale[n]) with E4M3 operands and fp32 accumulation (tcgen05
 * kind::f8f6f4), fp32 or bf16 output.  Replaces the torch._scaled_mm call sites of fp8_linear / fp8_fused_gate_up
 * (core/fp8.py:176-182, 236-242). */
int i64b_fp8_gemm(const uint8_t* x_q, const float* x_scale, const uint8_t* w_q, const float* w_scale, void* out,
                  int out_dtype, int64_t m, int64_t n, int64_t k, i64b_stream_t stream);

/* out_f32[m,n] = (x_q[m,:] . w_q[n,:]) * (x_scale[m] * w_scale[n])  — tcgen05 kind::i8, exact INT32.
 * Replaces int8_linear_native's torch._int_mm + rescale (core/quantization.py:119-176) and
 * triton_int8_gemm (kernels/triton/I64_int8_gemm.py:30-156).  acc_out (int32 [m,n]) optional. */
int i64b_int8_gemm(const int8_t* x_q, const float* x_scale, const int8_t* w_q, const float* w_scale,
                   float* out, int32_t* acc_out, int64_t m, int64_t n, int64_t k, i64b_stream_t stream);

/* out[m,n] = (sum_k a[m,k] * b[n,k]) * col_scale[n]   (bf16 inputs, fp32 accumulate; col_scale nullable;
 * out_dtype BF16 or F32).  With b = int8 weights held exactly in bf16 and col_scale = their per-channel scales
 * this replaces the W8A32 kernels I64_gemm_dequant_int8 / triton_dequant_gemm_int8 (csrc/I64_gemm.cu:40-78,
 * kernels/triton/I64_fused_dequant_gemm.py:32-146). */
int i64b_bf16_gemm(const void* a, const void* b, const float* col_scale, void* out, int out_dtype, int64_t m,
                   int64_t n, int64_t k, i64b_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* I64B_H_ */
