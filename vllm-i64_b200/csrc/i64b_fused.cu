// Fused decode-sized MoE FFN layer: ONE persistent launch (one CTA per SM) runs
//   plan (redundantly per CTA, N ≤ 1024) → row dispatch → grouped GEMM1 + SwiGLU → grouped GEMM2 → weighted combine
// with device-side dependencies instead of kernel boundaries.  Decode is bound by streaming the layer's weights from
// HBM; separate kernels pay prologue + first-byte latency + tail five times per layer and stop the stream at every
// boundary.  Here the TMA producer keeps prefetching WEIGHT tiles (which depend on nothing) across the
// dispatch→GEMM1 and GEMM1→GEMM2 dependencies; only the activation tiles wait on the global counters.
//
// Semantics are identical to the unfused chain (same plan, same tile math, same combine order), so the same parity
// tests cover both.  Reference functions covered: layers/token_routed_mlp.py:113-187 (route, expert_forward, forward).
#include <cuda.h>
#include "i64b_fused.cuh"

namespace {

constexpr int kThreads = 256;          // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue, warps 6-7 helpers
constexpr int kWarps = kThreads / 32;
constexpr int kBlockM = 128;
constexpr int kNMma = 128;             // UMMA N for both GEMMs (GEMM1: 64 gate + 64 up)
constexpr int kBn1 = 64;               // intermediate columns per GEMM1 tile
constexpr int kBlockK = 64;            // bf16 elements per 128-byte swizzle row
constexpr int kStageA = kBlockM * 128;
constexpr int kStageB = kNMma * 128;
constexpr int kStages = 5;
constexpr int kSlabK = 512;            // GEMM2 split-K slab (keeps every GEMM2 tile at ≤ 8 k-blocks)
constexpr uint32_t kTmemCols = 2 * kNMma;
constexpr int kMaxTok = I64B_FUSED_MAX_TOKENS;
constexpr int kMaxMTiles = I64B_FUSED_MAX_MTILES;

struct Smem {
    uint64_t full_bar[kStages], empty_bar[kStages], tfull_bar[2], tempty_bar[2];
    uint32_t tmem_ptr, pad;
    int group_rows[I64B_MAX_GROUPS];      // rows per group (experts, then shared)
    int group_start[I64B_MAX_GROUPS + 1]; // first xs / y row of each group
    int mt_prefix[I64B_MAX_GROUPS + 1];   // m-tiles before each group
    int t1_prefix[I64B_MAX_GROUPS + 1];   // GEMM1 tiles before each group
    int t2_prefix[I64B_MAX_GROUPS + 1];   // GEMM2 tiles before each group
    int route_base[I64B_MAX_BINS];
    int offsets[I64B_MAX_BINS];
    int warp_cnt[kWarps][I64B_MAX_BINS];
    uint16_t dest_row[I64B_MAX_TOPK][kMaxTok];
    uint8_t expert[I64B_MAX_TOPK][kMaxTok];
};

struct Tile {
    int type;      // 1 = GEMM1, 2 = GEMM2
    int g, m_t, n_t, slab;
    int nkb;
};

__device__ __forceinline__ uint64_t sw128_desc(const void* p) {
    uint64_t d = (uint64_t)((smem_u32(p) >> 4) & 0x3FFFu);
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= 1ull << 46;
    d |= 2ull << 61;
    return d;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float silu_f32(float g) { return __fdividef(g, 1.0f + __expf(-g)); }
__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ float4 ld_cg_f4(const float* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
// bounded spin on a device counter (a stuck dependency traps instead of hanging the GPU)
__device__ __forceinline__ void wait_counter(const int* ctr, int target) {
    const long long t0 = clock64();
    while (ld_acquire(ctr) < target) {
        if (clock64() - t0 > 4000000000ll) __trap();
    }
}

__device__ __forceinline__ void decode_tile(const I64bFusedParams& p, const Smem& s, int idx, int T1, Tile& t) {
    const int ng = p.num_experts + 1;
    if (idx < T1) {
        int g = 0;
        while (idx >= s.t1_prefix[g + 1]) ++g;
        const int n_out = g < p.num_experts ? p.expert_inter : p.shared_inter;
        const int nt = (n_out + kBn1 - 1) / kBn1;
        const int local = idx - s.t1_prefix[g];
        t.type = 1; t.g = g; t.m_t = local / nt; t.n_t = local % nt; t.slab = 0;
        t.nkb = (p.hidden + kBlockK - 1) / kBlockK;
    } else {
        const int j = idx - T1;
        int g = 0;
        while (j >= s.t2_prefix[g + 1]) ++g;
        const int kk = g < p.num_experts ? p.expert_inter : p.shared_inter;
        const int slabs = (kk + kSlabK - 1) / kSlabK;
        const int nt = (p.hidden + kNMma - 1) / kNMma;
        int local = j - s.t2_prefix[g];
        t.type = 2; t.g = g;
        t.slab = local % slabs; local /= slabs;
        t.n_t = local % nt; t.m_t = local / nt;
        const int krem = kk - t.slab * kSlabK;
        t.nkb = ((krem < kSlabK ? krem : kSlabK) + kBlockK - 1) / kBlockK;
    }
    (void)ng;
}

__global__ void __launch_bounds__(kThreads, 1)
i64b_fused_layer_kernel(const __grid_constant__ CUtensorMap tm_xs, const __grid_constant__ CUtensorMap tm_w13,
                        const __grid_constant__ CUtensorMap tm_inter_r, const __grid_constant__ CUtensorMap tm_inter_s,
                        const __grid_constant__ CUtensorMap tm_w2r, const __grid_constant__ CUtensorMap tm_w2s,
                        const __grid_constant__ I64bFusedParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = base;
    uint8_t* sB = base + kStages * kStageA;
    Smem& s = *reinterpret_cast<Smem*>(sB + kStages * kStageB);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int E = p.num_experts, K = p.top_k, N = p.n, H = p.hidden, bins = K * E, G = gridDim.x;
    pdl_launch_dependents();

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_xs); tma_prefetch_desc(&tm_w13); tma_prefetch_desc(&tm_inter_r);
        tma_prefetch_desc(&tm_inter_s); tma_prefetch_desc(&tm_w2r); tma_prefetch_desc(&tm_w2s);
        for (int i = 0; i < kStages; ++i) { mbar_init(&s.full_bar[i], 1); mbar_init(&s.empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s.tfull_bar[i], 1); mbar_init(&s.tempty_bar[i], 4); }
        fence_barrier_init();
    } else if (warp == 1) {
        tmem_alloc<kTmemCols>(&s.tmem_ptr);
    }
    for (int i = threadIdx.x; i < kWarps * I64B_MAX_BINS; i += kThreads) (&s.warp_cnt[0][0])[i] = 0;
    pdl_wait_prior();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = s.tmem_ptr;

    // ------------------------------------------------------------------ phase A: plan (every CTA, into smem)
    {
        const int tok_per_warp = ((N + kWarps - 1) / kWarps + 31) / 32 * 32;
        const int iters = tok_per_warp / 32;                       // ≤ 4 for N ≤ 1024
        const int tbase = warp * tok_per_warp;
        int e_reg[4][I64B_MAX_TOPK];
        uint16_t r_reg[4][I64B_MAX_TOPK];
        long long id[4];
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int t = tbase + it * 32 + lane;
            id[it] = (p.token_ids != nullptr && it < iters && t < N) ? p.token_ids[t] : 0;
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const long long c = id[it] < 0 ? 0 : (id[it] >= p.vocab ? p.vocab - 1 : id[it]);
#pragma unroll
            for (int k = 0; k < I64B_MAX_TOPK; ++k) {
                int e = 0;
                if (k < K) e = p.token_ids ? (int)__ldg(p.table + (int64_t)k * p.vocab + c) : (k % E);
                e_reg[it][k] = min(e, E - 1);
            }
        }
        const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            if (it < iters) {
                const int t = tbase + it * 32 + lane;
                const bool valid = t < N;
#pragma unroll
                for (int k = 0; k < I64B_MAX_TOPK; ++k) {
                    if (k < K) {
                        const int e = valid ? e_reg[it][k] : 0xFF;
                        const unsigned peers = __match_any_sync(0xffffffffu, e);
                        const int leader = __ffs(peers) - 1;
                        const int prev = valid ? s.warp_cnt[warp][k * E + e] : 0;
                        __syncwarp();
                        if (valid && lane == leader) s.warp_cnt[warp][k * E + e] = prev + __popc(peers);
                        __syncwarp();
                        r_reg[it][k] = (uint16_t)(prev + __popc(peers & lt_mask));
                    }
                }
            }
        }
        __syncthreads();
        for (int b = threadIdx.x; b < bins; b += kThreads) {       // exclusive scan over warps → per-bin totals
            int running = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const int c = s.warp_cnt[w][b];
                s.warp_cnt[w][b] = running;
                running += c;
            }
            s.offsets[b] = running;                                // holds counts until thread 0 scans them
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int gs = 0, mt = 0, t1 = 0, t2 = 0;
            const int nt2 = (H + kNMma - 1) / kNMma;
            for (int e = 0; e <= E; ++e) {
                int rows = 0;
                if (e < E) {
                    for (int k = 0; k < K; ++k) { s.route_base[k * E + e] = rows; rows += s.offsets[k * E + e]; }
                } else {
                    rows = p.shared_inter > 0 ? N : 0;
                }
                const int n_out = e < E ? p.expert_inter : p.shared_inter;
                const int kk = n_out;
                s.group_rows[e] = rows; s.group_start[e] = gs; s.mt_prefix[e] = mt; s.t1_prefix[e] = t1; s.t2_prefix[e] = t2;
                const int m = (rows + kBlockM - 1) / kBlockM;
                mt += m;
                t1 += m * ((n_out + kBn1 - 1) / kBn1);
                t2 += m * nt2 * ((kk + kSlabK - 1) / kSlabK);
                gs += rows;
            }
            s.group_start[E + 1] = gs; s.mt_prefix[E + 1] = mt; s.t1_prefix[E + 1] = t1; s.t2_prefix[E + 1] = t2;
            if (blockIdx.x == 0 && p.tap_counts) {
                for (int b = 0; b < bins; ++b) p.tap_counts[b] = s.offsets[b];
            }
            for (int k = 0; k < K; ++k) {                          // per-route exclusive offsets (reference convention)
                int off = 0;
                for (int e = 0; e < E; ++e) { const int c = s.offsets[k * E + e]; s.offsets[k * E + e] = off; off += c; }
            }
            if (blockIdx.x == 0 && p.tap_offsets)
                for (int b = 0; b < bins; ++b) p.tap_offsets[b] = s.offsets[b];
        }
        __syncthreads();
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int t = tbase + it * 32 + lane;
            if (it < iters && t < N) {
#pragma unroll
                for (int k = 0; k < I64B_MAX_TOPK; ++k)
                    if (k < K) {
                        const int e = e_reg[it][k];
                        const int bin = k * E + e;
                        const int in_exp = s.warp_cnt[warp][bin] + r_reg[it][k];
                        const int dest = s.group_start[e] + s.route_base[bin] + in_exp;
                        s.dest_row[k][t] = (uint16_t)dest;
                        s.expert[k][t] = (uint8_t)e;
                        if (blockIdx.x == 0) {
                            if (p.tap_routes) p.tap_routes[(int64_t)k * N + t] = e;
                            if (p.tap_dest) p.tap_dest[(int64_t)k * N + t] = dest;
                            if (p.tap_scatter) p.tap_scatter[(int64_t)k * N + t] = s.offsets[bin] + in_exp;
                        }
                    }
            }
        }
        __syncthreads();
    }
    const int R = s.group_start[E];                  // routed rows (= N*K)
    const int T1 = s.t1_prefix[E + 1], T2 = s.t2_prefix[E + 1];
    const int total_tiles = T1 + T2;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int b_idx = blockIdx.x, a_idx = blockIdx.x;      // tile cursors for the weight and the activation loads
            int b_kb = 0, a_kb = 0, b_stage = 0, a_stage = 0, issued_ahead = 0;
            uint32_t b_phase = 0;
            Tile bt, at;
            bool bt_valid = false, at_valid = false, at_ready = false;
            const long long t_begin = clock64();
            while (a_idx < total_tiles) {
                // weights: run ahead as far as the ring allows
                while (b_idx < total_tiles && issued_ahead < kStages) {
                    if (!bt_valid) { decode_tile(p, s, b_idx, T1, bt); bt_valid = true; }
                    if (!mbar_try_wait(&s.empty_bar[b_stage], b_phase ^ 1)) break;
                    mbar_arrive_expect_tx(&s.full_bar[b_stage], kStageA + kStageB);
                    uint8_t* dstB = sB + b_stage * kStageB;
                    if (bt.type == 1) {
                        const int n_out = bt.g < E ? p.expert_inter : p.shared_inter;
                        const int wrow = (bt.g < E ? bt.g * 2 * p.expert_inter : E * 2 * p.expert_inter) + bt.n_t * kBn1;
                        tma_load_2d(dstB, &tm_w13, &s.full_bar[b_stage], b_kb * kBlockK, wrow, I64B_L2_EVICT_FIRST);
                        tma_load_2d(dstB + kBn1 * 128, &tm_w13, &s.full_bar[b_stage], b_kb * kBlockK, wrow + n_out,
                                    I64B_L2_EVICT_FIRST);
                    } else {
                        const CUtensorMap* tb = bt.g < E ? &tm_w2r : &tm_w2s;
                        const int wrow = (bt.g < E ? bt.g * H : 0) + bt.n_t * kNMma;
                        tma_load_2d(dstB, tb, &s.full_bar[b_stage], bt.slab * kSlabK + b_kb * kBlockK, wrow,
                                    I64B_L2_EVICT_FIRST);
                    }
                    ++issued_ahead;
                    if (++b_stage == kStages) { b_stage = 0; b_phase ^= 1; }
                    if (++b_kb == bt.nkb) { b_kb = 0; b_idx += G; bt_valid = false; }
                }
                // activations: need the producing phase of other CTAs
                if (issued_ahead > 0) {
                    if (!at_valid) { decode_tile(p, s, a_idx, T1, at); at_valid = true; at_ready = false; }
                    if (!at_ready) {
                        const int* ctr = at.type == 1 ? p.sync + I64B_SYNC_DISPATCH
                                                      : p.sync + I64B_SYNC_G1 + s.mt_prefix[at.g] + at.m_t;
                        const int n_out = at.g < E ? p.expert_inter : p.shared_inter;
                        const int target = at.type == 1 ? N : (n_out + kBn1 - 1) / kBn1;
                        if (ld_acquire(ctr) >= target) { fence_proxy_async_global(); at_ready = true; }
                        else if (clock64() - t_begin > 4000000000ll) __trap();
                    }
                    if (at_ready) {
                        uint8_t* dstA = sA + a_stage * kStageA;
                        const int arow = at.m_t * kBlockM;
                        if (at.type == 1) {
                            tma_load_2d(dstA, &tm_xs, &s.full_bar[a_stage], a_kb * kBlockK, s.group_start[at.g] + arow,
                                        I64B_L2_EVICT_NORMAL);
                        } else if (at.g < E) {
                            tma_load_2d(dstA, &tm_inter_r, &s.full_bar[a_stage], at.slab * kSlabK + a_kb * kBlockK,
                                        s.group_start[at.g] + arow, I64B_L2_EVICT_NORMAL);
                        } else {
                            tma_load_2d(dstA, &tm_inter_s, &s.full_bar[a_stage], at.slab * kSlabK + a_kb * kBlockK, arow,
                                        I64B_L2_EVICT_NORMAL);
                        }
                        --issued_ahead;
                        if (++a_stage == kStages) a_stage = 0;
                        if (++a_kb == at.nkb) { a_kb = 0; a_idx += G; at_valid = false; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            // c_format F32 (1<<4), A/B bf16 (1<<7, 1<<10), N>>3 at 17, M>>4 at 24
            constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kNMma >> 3) << 17) |
                                       ((uint32_t)(kBlockM >> 4) << 24);
            int stage = 0, iter = 0;
            uint32_t phase = 0;
            for (int idx = blockIdx.x; idx < total_tiles; idx += G, ++iter) {
                Tile t;
                decode_tile(p, s, idx, T1, t);
                const int as = iter & 1;
                mbar_wait(&s.tempty_bar[as], ((iter >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * kNMma;
                for (int kb = 0; kb < t.nkb; ++kb) {
                    mbar_wait(&s.full_bar[stage], phase);
                    tc_fence_after();
                    const uint64_t a_desc = sw128_desc(sA + stage * kStageA);
                    const uint64_t b_desc = sw128_desc(sB + stage * kStageB);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                    umma_commit(&s.empty_bar[stage]);
                    if (kb == t.nkb - 1) umma_commit(&s.tfull_bar[as]);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp < 6) {
        // ------------------------------------------------------------------ epilogue warps
        const int q = warp & 3;
        const int r = q * 32 + lane;
        int iter = 0;
        for (int idx = blockIdx.x; idx < total_tiles; idx += G, ++iter) {
            Tile t;
            decode_tile(p, s, idx, T1, t);
            const int as = iter & 1;
            mbar_wait(&s.tfull_bar[as], (iter >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * kNMma;
            const int row_in_group = t.m_t * kBlockM + r;
            const bool valid = row_in_group < s.group_rows[t.g];
            if (t.type == 1) {
                const bool shared = t.g >= E;
                const int n_out = shared ? p.shared_inter : p.expert_inter;
                __nv_bfloat16* orow = shared
                    ? reinterpret_cast<__nv_bfloat16*>(p.inter_s) + (int64_t)row_in_group * p.shared_inter
                    : reinterpret_cast<__nv_bfloat16*>(p.inter_r) + (int64_t)(s.group_start[t.g] + row_in_group) * p.expert_inter;
#pragma unroll 1
                for (int c = 0; c < kBn1 / 32; ++c) {
                    const int col0 = t.n_t * kBn1 + c * 32;
                    if (col0 >= n_out) break;
                    uint32_t gr[32], ur[32];
                    tmem_ld_32x32(taddr + c * 32, gr);
                    tmem_ld_32x32(taddr + kBn1 + c * 32, ur);
                    tmem_ld_wait();
                    uint32_t o[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        o[j] = pack_bf16x2(silu_f32(__uint_as_float(gr[2 * j])) * __uint_as_float(ur[2 * j]),
                                           silu_f32(__uint_as_float(gr[2 * j + 1])) * __uint_as_float(ur[2 * j + 1]));
                    if (valid) {
#pragma unroll
                        for (int v = 0; v < 4; ++v)
                            if (col0 + v * 8 < n_out)
                                *reinterpret_cast<int4*>(orow + col0 + v * 8) =
                                    make_int4(o[4 * v], o[4 * v + 1], o[4 * v + 2], o[4 * v + 3]);
                    }
                }
            } else {
                const bool shared = t.g >= E;
                const int64_t yrow = shared ? (int64_t)R + (int64_t)t.slab * N + row_in_group
                                            : (int64_t)s.group_start[t.g] + row_in_group;
                float* orow = p.y + yrow * H;
#pragma unroll 1
                for (int c = 0; c < kNMma / 32; ++c) {
                    const int col0 = t.n_t * kNMma + c * 32;
                    if (col0 >= H) break;
                    uint32_t ar[32];
                    tmem_ld_32x32(taddr + c * 32, ar);
                    tmem_ld_wait();
                    if (valid) {
#pragma unroll
                        for (int v = 0; v < 8; ++v)
                            if (col0 + v * 4 < H)
                                *reinterpret_cast<int4*>(orow + col0 + v * 4) =
                                    make_int4(ar[4 * v], ar[4 * v + 1], ar[4 * v + 2], ar[4 * v + 3]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s.tempty_bar[as]);
            // publish the tile: all four epilogue warps' stores → one release increment
            __threadfence();
            named_bar_sync(1, 128);
            if (warp == 2 && lane == 0) {
                if (t.type == 1) atomicAdd(p.sync + I64B_SYNC_G1 + s.mt_prefix[t.g] + t.m_t, 1);
                else atomicAdd(p.sync + I64B_SYNC_G2, 1);
            }
        }
    }

    // ------------------------------------------------------------------ phase B: dispatch rows (warps 6,7 start at
    // once; the role warps join when their tile loops are empty — with N ≤ 1024 each CTA copies at most a few rows)
    if (warp >= 6) {
        const int nvec = H / 8;
        int copied = 0;
        for (int t = blockIdx.x + G * (warp - 6); t < N; t += 2 * G) {
            const int4* src = reinterpret_cast<const int4*>(reinterpret_cast<const __nv_bfloat16*>(p.x) + (int64_t)t * H);
            for (int i = lane; i < nvec; i += 32) {
                const int4 v = ld_nc_16(src + i);
                for (int k = 0; k < K; ++k)
                    *reinterpret_cast<int4*>(reinterpret_cast<__nv_bfloat16*>(p.xs) + (int64_t)s.dest_row[k][t] * H + i * 8) = v;
                if (p.shared_inter > 0)
                    *reinterpret_cast<int4*>(reinterpret_cast<__nv_bfloat16*>(p.xs) + (int64_t)(R + t) * H + i * 8) = v;
            }
            ++copied;
        }
        __threadfence();
        __syncwarp();
        if (lane == 0 && copied) atomicAdd(p.sync + I64B_SYNC_DISPATCH, copied);
    }

    // ------------------------------------------------------------------ phase E: combine this CTA's tokens
    __syncthreads();
    if (threadIdx.x == 0) wait_counter(p.sync + I64B_SYNC_G2, T2);
    __syncthreads();
    {
        const int hv = H / 4;
        const int slabs_s = p.shared_inter > 0 ? (p.shared_inter + kSlabK - 1) / kSlabK : 0;
        for (int t = blockIdx.x; t < N; t += G) {
            for (int c = threadIdx.x; c < hv; c += kThreads) {
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                if (E > 0) {
#pragma unroll
                    for (int k = 0; k < I64B_MAX_TOPK; ++k) {
                        if (k < K) {
                            const float4 y = ld_cg_f4(p.y + (int64_t)s.dest_row[k][t] * H + c * 4);
                            const float w = p.w[k];
                            if (k == 0) {
                                o = make_float4(__fmul_rn(w, y.x), __fmul_rn(w, y.y), __fmul_rn(w, y.z), __fmul_rn(w, y.w));
                            } else {
                                o.x = __fadd_rn(o.x, __fmul_rn(w, y.x)); o.y = __fadd_rn(o.y, __fmul_rn(w, y.y));
                                o.z = __fadd_rn(o.z, __fmul_rn(w, y.z)); o.w = __fadd_rn(o.w, __fmul_rn(w, y.w));
                            }
                        }
                    }
                    const float rs = p.routed_scale;
                    o = make_float4(__fmul_rn(rs, o.x), __fmul_rn(rs, o.y), __fmul_rn(rs, o.z), __fmul_rn(rs, o.w));
                }
                if (slabs_s > 0) {
                    float4 sh = ld_cg_f4(p.y + ((int64_t)R + t) * H + c * 4);
                    for (int sl = 1; sl < slabs_s; ++sl) {      // split-K slabs of the shared down-projection, fixed order
                        const float4 y = ld_cg_f4(p.y + ((int64_t)R + (int64_t)sl * N + t) * H + c * 4);
                        sh.x = __fadd_rn(sh.x, y.x); sh.y = __fadd_rn(sh.y, y.y);
                        sh.z = __fadd_rn(sh.z, y.z); sh.w = __fadd_rn(sh.w, y.w);
                    }
                    if (E > 0) {
                        const float ss = p.shared_scale;
                        o.x = __fadd_rn(o.x, __fmul_rn(ss, sh.x)); o.y = __fadd_rn(o.y, __fmul_rn(ss, sh.y));
                        o.z = __fadd_rn(o.z, __fmul_rn(ss, sh.z)); o.w = __fadd_rn(o.w, __fmul_rn(ss, sh.w));
                    } else {
                        o = sh;
                    }
                }
                if (p.out_dtype == I64B_F32) {
                    reinterpret_cast<float4*>(p.out)[(int64_t)t * hv + c] = o;
                } else {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
                    reinterpret_cast<uint2*>(p.out)[(int64_t)t * hv + c] =
                        make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
                }
            }
        }
    }

    // ------------------------------------------------------------------ teardown: last CTA resets the counters
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
    if (threadIdx.x == 0) {
        __threadfence();
        const int ticket = atomicAdd(p.sync + I64B_SYNC_FINISH, 1);
        if (ticket == G - 1) {
            for (int i = 0; i < I64B_SYNC_INTS; ++i) p.sync[i] = 0;
            __threadfence();
        }
    }
}

constexpr int kSmemBytes = 1024 + kStages * (kStageA + kStageB) + (int)sizeof(Smem) + 256;

}  // namespace

int i64b_fused_supported(int n, int hidden, int expert_inter, int shared_inter, int num_experts, int top_k) {
    if (n < 1 || n > kMaxTok) return 0;
    if (hidden % 64 != 0 || hidden > 8192) return 0;
    if (num_experts > 0 && expert_inter % 16 != 0) return 0;
    if (shared_inter % 16 != 0) return 0;
    const int64_t mt = (int64_t)n * top_k / kBlockM + num_experts + (n + kBlockM - 1) / kBlockM;
    if (mt > kMaxMTiles) return 0;
    if ((int64_t)n * (top_k + 1) >= 65536) return 0;       // uint16 destination rows
    static_assert(kSmemBytes <= 232448, "shared memory budget");
    return 1;
}

int i64b_launch_fused(const CUtensorMap& tm_xs, const CUtensorMap& tm_w13, const CUtensorMap& tm_inter_r,
                      const CUtensorMap& tm_inter_s, const CUtensorMap& tm_w2r, const CUtensorMap& tm_w2s,
                      const I64bFusedParams& p, cudaStream_t stream) {
    static bool configured = false;
    if (!configured) {
        I64B_CUDA(cudaFuncSetAttribute(i64b_fused_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        configured = true;
    }
    const int grid = i64b_num_sms();
    I64B_CUDA(i64b_launch(i64b_fused_layer_kernel, dim3(grid), dim3(kThreads), kSmemBytes, stream, tm_xs, tm_w13,
                          tm_inter_r, tm_inter_s, tm_w2r, tm_w2s, p));
    i64b_count_launch(1);
    return I64B_OK;
}
