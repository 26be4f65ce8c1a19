// Fused decode-layer kernel interface (see i64b_fused.cu).
#pragma once
#include <cuda.h>
#include "i64b_common.cuh"

#define I64B_FUSED_MAX_TOKENS 1024
#define I64B_FUSED_MAX_GROUPS 128     // experts + shared m-tiles (one GEMM1→GEMM2 counter each)
// persistent per-layer sync area (ints, zero at load, reset by the last CTA of every launch)
// every counter sits on its own 128-byte line: pollers of one dependency do not queue behind another's at L2
#define I64B_SYNC_STRIDE 32
#define I64B_SYNC_G2 (1 * I64B_SYNC_STRIDE)
#define I64B_SYNC_FINISH (2 * I64B_SYNC_STRIDE)
#define I64B_SYNC_G1(i) ((4 + (i)) * I64B_SYNC_STRIDE)
#define I64B_SYNC_INTS ((4 + I64B_FUSED_MAX_GROUPS) * I64B_SYNC_STRIDE)

// static work list of a launch (a pure function of the shapes; computed on the host)
struct I64bFusedGeo {
    int nkb_h;                 // k-blocks over H
    int nt1r, U1r;             // GEMM1 routed: n-tiles per expert, units
    int mts, nt1s, U1;         // shared m-tiles, GEMM1 shared n-tiles, total GEMM1 units
    int nt2, U2r, slabs, U2;   // GEMM2 n-tiles, routed units, shared K slabs, total GEMM2 units
    int U;
    unsigned m_nt1r, m_nt1s, m_nt2, m_slabs;   // ceil(2^32 / d) of the four divisors of the unit decode (0 for d <= 1)
};

struct I64bFusedParams {
    const void* x;              // [N, H] bf16
    const int64_t* token_ids;   // [N] or nullptr
    const uint8_t* table;       // [K, V] uint8
    int32_t n, hidden, num_experts, expert_inter, shared_inter, top_k, vocab, out_dtype;
    float w[I64B_MAX_TOPK];
    float routed_scale, shared_scale;
    void* inter_r;              // [N*K, I_e] bf16, expert-major rows
    void* inter_s;              // [N, S] bf16
    float* y;                   // [N*K + slabs*N, H] fp32
    void* out;                  // [N, H]
    int* sync;                  // I64B_SYNC_INTS ints
    const uint8_t* next_w;      // next layer's packed weights (constants) to pull into L2 while this layer runs, or nullptr
    uint64_t next_w_bytes;
    int32_t prefetch_mode;
    I64bFusedGeo geo;           // filled by i64b_launch_fused      // 1: a spare thread at kernel entry; 2: the weight thread once its own stream is issued
    long long* timeline;        // bring-up only: [grid, 16] ns stamps (env I64B_DEBUG_TIMELINE)
    int32_t *tap_routes, *tap_dest, *tap_scatter, *tap_counts, *tap_offsets;   // reference-convention plan outputs
};

int i64b_fused_supported(int n, int hidden, int expert_inter, int shared_inter, int num_experts, int top_k);
int i64b_launch_fused(const CUtensorMap& tm_x, const CUtensorMap& tm_w13, const CUtensorMap& tm_inter_r, const CUtensorMap& tm_inter_s,
                      const CUtensorMap& tm_w2r, const CUtensorMap& tm_w2s, const I64bFusedParams& p,
                      cudaStream_t stream);
int i64b_num_sms();
void i64b_count_launch(int n);
