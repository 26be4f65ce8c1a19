// Fused decode-sized MoE FFN layer: ONE persistent launch (one CTA per SM) runs
//   route + plan (per CTA, in shared memory) → grouped GEMM1 + SwiGLU → grouped GEMM2 → weighted combine
// with device-side dependencies instead of kernel boundaries.
//
// Decode is bound by streaming the layer's weights from HBM exactly once, so the kernel is organised around the
// WEIGHT stream:
//   * the work list is STATIC (independent of the routing): unit = one 128-row weight tile of one expert (or of the
//     shared expert) over its whole K range.  Every CTA therefore knows at entry which weight bytes it will need and
//     its TMA thread starts pulling them before the plan exists, before the previous kernel has finished (weights
//     are constants: the loads sit above griddepcontrol.wait) and across the GEMM1→GEMM2 dependency;
//   * there is no dispatch pass and no dispatched-row buffer: four gather warps copy the A rows of a routed GEMM1
//     tile straight from x into the 128B-swizzled stage (cp.async, 16 B per lane, four full 128-byte lines per warp
//     instruction, completion signalled on the stage's mbarrier by cp.async.mbarrier.arrive) using the plan's
//     expert-sorted token list, so GEMM1 depends only on the CTA-local plan (TMA tile::gather4 was measured at
//     1.5x the time per k-block: 512-byte requests are too small for the TMA unit);
//   * the only grid-wide dependencies left are GEMM1→GEMM2 (per expert / per shared m-tile counters) and
//     GEMM2→combine (one counter); the last CTA to leave resets them, so back-to-back launches need no memset.
// An expert that owns more than 128 rows is processed in passes of 128 rows; its weight tile is streamed again for
// every pass (from L2).  Semantics (plan, tile math, combine order) match the unfused chain; the same parity tests
// cover both.  Reference functions covered: layers/token_routed_mlp.py:113-187 (route, expert_forward, forward).
#include <cuda.h>
#include <stdlib.h>
#include "i64b_fused.cuh"

namespace {

constexpr int kThreads = 352;          // warp 0 weight TMA, 1 MMA, 2-5 epilogue, 6-9 A-row gather, 10 activation TMA
constexpr int kPlanWarps = 8;          // warps 2-9 build the plan before taking their roles
constexpr int kGatherThreads = 128;
constexpr int kBlockM = 128;
constexpr int kNMma = 128;             // UMMA N for both GEMMs (GEMM1: 64 gate + 64 up)
constexpr int kBn1 = 64;               // intermediate columns per GEMM1 unit
constexpr int kBlockK = 64;            // bf16 elements per 128-byte swizzle row
constexpr int kStageA = kBlockM * 128;
constexpr int kStageB = kNMma * 128;
constexpr int kStages = 5;
constexpr int kStgPitch = 144;       // bytes per staged row: 128 + 16 (conflict-free 16-byte writes by row, reads by line)
constexpr int kStgBytes = 4 * 32 * kStgPitch;   // one 32-row x 128-byte staging tile per epilogue warp
constexpr int kMaxSlabs = 8;          // split-K slabs of the shared down-projection (shared_inter <= 4096)
constexpr int kSlabK = 512;            // split-K slab of the shared down-projection
constexpr uint32_t kTmemCols = 2 * kNMma;
constexpr int kMaxTok = I64B_FUSED_MAX_TOKENS;

constexpr int kMaxUnitsPerCta = 8;
enum { kG1Routed = 0, kG1Shared = 1, kG2Routed = 2, kG2Shared = 3 };
struct Unit {
    int type, g, n_t, m_t, slab, nkb;
};

struct Smem {
    uint64_t full_bar[kStages], empty_bar[kStages], tfull_bar[2], tempty_bar[2];
    uint32_t tmem_ptr;
    uint32_t b_seq;                          // weight stages issued so far (release/acquire, CTA scope)
    int group_rows[I64B_MAX_EXPERTS];        // routed rows per expert
    int group_start[I64B_MAX_EXPERTS + 1];   // first inter_r / y row of each expert; [E] = N*K
    int route_base[I64B_MAX_BINS];
    int offsets[I64B_MAX_BINS];
    uint16_t dest_row[I64B_MAX_TOPK][kMaxTok];          // merged expert-major row of assignment (k, t)
    uint16_t src_tok[I64B_MAX_TOPK * kMaxTok];          // inverse: token of merged row
    uint8_t expert[I64B_MAX_TOPK][kMaxTok];             // 0xFF past N up to the next multiple of 128
    // fast path for the CTA's first unit (routed GEMM1): the first 128 tokens of its expert, and the expert's row count
    uint16_t own_tok[kBlockM];
    int fast_cnt[kPlanWarps];
    int rows_first;
    uint32_t roles_go;                                  // everybody but the weight warp is past the role barrier
    uint32_t plan_ready;                                // full plan published (release/acquire, CTA scope)
    Unit units[kMaxUnitsPerCta];                        // this CTA's units (static work list), decoded once by warp 1
    int n_units;
    int comb_rank, comb_total;                          // this CTA's index among the combine workers (-1: none) and their number
};

// static work list
using Geo = I64bFusedGeo;

static Geo make_geo(const I64bFusedParams& p) {
    Geo g;
    const int E = p.num_experts;
    g.nkb_h = (p.hidden + kBlockK - 1) / kBlockK;
    g.nt1r = E > 0 ? (p.expert_inter + kBn1 - 1) / kBn1 : 0;
    g.U1r = E * g.nt1r;
    g.mts = p.shared_inter > 0 ? (p.n + kBlockM - 1) / kBlockM : 0;
    g.nt1s = p.shared_inter > 0 ? (p.shared_inter + kBn1 - 1) / kBn1 : 0;
    g.U1 = g.U1r + g.mts * g.nt1s;
    g.nt2 = (p.hidden + kNMma - 1) / kNMma;
    g.U2r = E * g.nt2;
    g.slabs = p.shared_inter > 0 ? (p.shared_inter + kSlabK - 1) / kSlabK : 0;
    g.U2 = g.U2r + g.mts * g.nt2 * g.slabs;
    g.U = g.U1 + g.U2;
    auto magic = [](int d) { return d <= 1 ? 0u : (unsigned)((0x100000000ull + (unsigned)d - 1) / (unsigned)d); };
    g.m_nt1r = magic(g.nt1r); g.m_nt1s = magic(g.nt1s); g.m_nt2 = magic(g.nt2); g.m_slabs = magic(g.slabs);
    return g;
}

// v / d for v * d < 2^32 with m = ceil(2^32 / d) (exact: the unit indices are small); d <= 1 has m == 0
__device__ __forceinline__ int fast_div(int v, unsigned m) { return m == 0 ? v : (int)__umulhi((unsigned)v, m); }
__device__ __noinline__ void decode_unit(const I64bFusedParams& p, const Geo& geo, int v, Unit& u) {
    u.m_t = 0; u.slab = 0;
    if (v < geo.U1r) {
        u.type = kG1Routed; u.g = fast_div(v, geo.m_nt1r); u.n_t = v - u.g * geo.nt1r; u.nkb = geo.nkb_h;
    } else if (v < geo.U1) {
        const int w = v - geo.U1r;
        u.type = kG1Shared; u.g = p.num_experts; u.m_t = fast_div(w, geo.m_nt1s); u.n_t = w - u.m_t * geo.nt1s; u.nkb = geo.nkb_h;
    } else if (v < geo.U1 + geo.U2r) {
        const int w = v - geo.U1;
        u.type = kG2Routed; u.g = fast_div(w, geo.m_nt2); u.n_t = w - u.g * geo.nt2;
        u.nkb = (p.expert_inter + kBlockK - 1) / kBlockK;
    } else {
        int w = v - geo.U1 - geo.U2r;
        u.type = kG2Shared; u.g = p.num_experts;
        const int q = fast_div(w, geo.m_slabs);
        u.slab = w - q * geo.slabs;
        u.m_t = fast_div(q, geo.m_nt2); u.n_t = q - u.m_t * geo.nt2;
        const int krem = p.shared_inter - u.slab * kSlabK;
        u.nkb = ((krem < kSlabK ? krem : kSlabK) + kBlockK - 1) / kBlockK;
    }
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
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
__device__ __forceinline__ void red_release_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_n(uint64_t* bar, uint32_t n) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}
// one arrival on `bar` once every cp.async this thread has issued so far has landed
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void st_release_cta(uint32_t* p, uint32_t v) {
    asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" :: "r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_cta(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
// The weight thread is the only one that observes every phase of the empty barriers; the activation roles skip the
// stages of units that are not theirs, so they may not wait on a parity.  They wait until the weight tile of stage
// `seq` has been issued instead, which the weight thread does right after the slot was released.
__device__ __forceinline__ void wait_stage_owned(const uint32_t* b_seq, uint32_t seq) {
    if (ld_acquire_cta(b_seq) > seq) return;
    const long long t0 = clock64();
    while (ld_acquire_cta(b_seq) <= seq) {
        __nanosleep(40);                  // four gather warps spin here: keep them off the issue slots of the epilogue warps
        if (clock64() - t0 > 4000000000ll) __trap();
    }
}
__device__ __noinline__ void wait_plan_slow(const uint32_t* flag) {
    const long long t0 = clock64();
    while (ld_acquire_cta(flag) == 0)
        if (clock64() - t0 > 4000000000ll) __trap();
}
// Rows of a routed unit's expert (-1 for the shared expert).  The CTA's first unit gets its count from the fast path;
// everything else needs the full plan, which the epilogue warps publish a little later.
__device__ __forceinline__ int unit_rows(const Smem& s, const Unit& u, bool first, bool& have_plan) {
    if (u.type == kG1Shared || u.type == kG2Shared) return -1;
    if (first && u.type == kG1Routed) return s.rows_first;
    if (!have_plan) {
        if (ld_acquire_cta(&s.plan_ready) == 0) wait_plan_slow(&s.plan_ready);
        have_plan = true;
    }
    return s.group_rows[u.g];
}
// 128-row passes over a unit's rows (an expert without tokens still runs one, empty, pass: its weights are already
// on their way when the plan completes)
__device__ __forceinline__ int passes_of(int rows) { return rows <= kBlockM ? 1 : (rows + kBlockM - 1) / kBlockM; }
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
    asm volatile("bar.arrive %0, %1;" :: "r"(id), "r"(threads) : "memory");
}
// mbarrier wait that traps instead of hanging the GPU when a dependency never arrives
__device__ __noinline__ void mbar_wait_slow(uint64_t* bar, uint32_t parity) {
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000ll) __trap();
    }
}
__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    if (!mbar_try_wait(bar, parity)) mbar_wait_slow(bar, parity);
}
__device__ __forceinline__ float4 ld_cg_f4(const float* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return (long long)t;
}
// Cross-layer weight prefetch: HBM is idle for most of a decode layer (the time goes to dependency latency), so each
// launch pulls its share of the NEXT layer's weights (constants) into the 126 MB L2; that layer's TMA stream then
// hits L2 instead of paying loaded-DRAM latency per ring refill.
__device__ __noinline__ void prefetch_next_weights(const I64bFusedParams& p, int cta, int ctas) {
    constexpr uint64_t kChunk = 16384;
    const uint64_t chunks = (p.next_w_bytes + kChunk - 1) / kChunk;
    for (uint64_t c = (uint64_t)cta; c < chunks; c += (uint64_t)ctas) {
        const uint64_t off = c * kChunk;
        const uint32_t bytes = (uint32_t)((p.next_w_bytes - off < kChunk ? p.next_w_bytes - off : kChunk) & ~uint64_t(15));
        if (bytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p.next_w + off), "r"(bytes) : "memory");
    }
}
#define TL(slot) do { if (kTL && tl) tl[slot] = gtime(); } while (0)
// bounded spin on a device counter (a stuck dependency traps instead of hanging the GPU)
__device__ __forceinline__ void wait_counter(const int* ctr, int target) {
    const long long t0 = clock64();
    while (ld_acquire(ctr) < target) {
        __nanosleep(64);
        if (clock64() - t0 > 4000000000ll) __trap();
    }
}

__device__ __noinline__ void issue_weights(const I64bFusedParams& p, const CUtensorMap* tm_w13, const CUtensorMap* tm_w2r,
                                           const CUtensorMap* tm_w2s, uint8_t* dstB, uint64_t* bar, int type, int g, int n_t,
                                           int slab, int kb) {
    mbar_arrive_expect_tx(bar, kStageB);
    if (type <= kG1Shared) {
        const int n_out = type == kG1Routed ? p.expert_inter : p.shared_inter;
        const int wrow = (type == kG1Routed ? g * 2 * p.expert_inter : p.num_experts * 2 * p.expert_inter) + n_t * kBn1;
        tma_load_2d(dstB, tm_w13, bar, kb * kBlockK, wrow, I64B_L2_EVICT_FIRST);
        tma_load_2d(dstB + kBn1 * 128, tm_w13, bar, kb * kBlockK, wrow + n_out, I64B_L2_EVICT_FIRST);
    } else if (type == kG2Routed) {
        tma_load_2d(dstB, tm_w2r, bar, kb * kBlockK, g * p.hidden + n_t * kNMma, I64B_L2_EVICT_FIRST);
    } else {
        tma_load_2d(dstB, tm_w2s, bar, slab * kSlabK + kb * kBlockK, n_t * kNMma, I64B_L2_EVICT_FIRST);
    }
}

template <bool kTL>      // kTL: bring-up time stamps compiled in (a separate instantiation: the product kernel stays small)
__global__ void __launch_bounds__(kThreads, 1)
i64b_fused_layer_kernel(const __grid_constant__ CUtensorMap tm_x,
                        const __grid_constant__ CUtensorMap tm_w13, const __grid_constant__ CUtensorMap tm_inter_r,
                        const __grid_constant__ CUtensorMap tm_inter_s, const __grid_constant__ CUtensorMap tm_w2r,
                        const __grid_constant__ CUtensorMap tm_w2s, const __grid_constant__ I64bFusedParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = base;
    uint8_t* sB = base + kStages * kStageA;
    uint8_t* sStg = sB + kStages * kStageB;
    Smem& s = *reinterpret_cast<Smem*>(sStg + kStgBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int E = p.num_experts, K = p.top_k, N = p.n, H = p.hidden, bins = K * E, G = gridDim.x;
    const Geo& geo = p.geo;
    pdl_launch_dependents();
    long long* tl = (kTL && p.timeline) ? p.timeline + (size_t)blockIdx.x * 32 : nullptr;
    if (threadIdx.x == 0) TL(0);

    // weight tile of (unit, k-block) → shared-memory stage; tx bytes = kStageB (one out-of-line copy: every TMA call
    // site costs ~17 instructions of uniform-register plumbing, and this kernel runs its code cold)
    auto issue_b = [&](const Unit& u, int kb, int stage) {
        issue_weights(p, &tm_w13, &tm_w2r, &tm_w2s, sB + stage * kStageB, &s.full_bar[stage], u.type, u.g, u.n_t, u.slab, kb);
    };

    int own_e = -1;              // expert of the CTA's first unit when that is a routed GEMM1 unit
    if ((int)blockIdx.x < geo.U1r) own_e = fast_div((int)blockIdx.x, geo.m_nt1r);
    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&tm_w13); tma_prefetch_desc(&tm_w2r); tma_prefetch_desc(&tm_w2s);
            tma_prefetch_desc(&tm_inter_r); tma_prefetch_desc(&tm_inter_s);
            tma_prefetch_desc(&tm_x);
            for (int i = 0; i < kStages; ++i) { mbar_init(&s.full_bar[i], 1 + kGatherThreads); mbar_init(&s.empty_bar[i], 1); }
            for (int i = 0; i < 2; ++i) { mbar_init(&s.tfull_bar[i], 1); mbar_init(&s.tempty_bar[i], 4); }
            fence_barrier_init();
            s.b_seq = 0;
            s.plan_ready = 0;
            s.roles_go = 0;
            __threadfence_block();
        }
        __syncwarp();
        // The weight warp only ARRIVES at the role barrier (its barriers and counters are ready) and goes on issuing:
        // it used to be the last one there (0.25-0.5 us per stage issued), holding back everybody else's start.
        named_bar_arrive(3, kThreads);
    } else if (warp == 1) {
        tmem_alloc<kTmemCols>(&s.tmem_ptr);
        // this CTA's units, decoded once (the role loops below read them from shared memory)
        if (lane < kMaxUnitsPerCta && (int)blockIdx.x + lane * G < geo.U) decode_unit(p, geo, blockIdx.x + lane * G, s.units[lane]);
        if (lane == 0) s.n_units = (int)blockIdx.x < geo.U ? (geo.U - (int)blockIdx.x + G - 1) / G : 0;
        // Combine: token t belongs to CTA t % G.  (Measured: handing the combine to the ~36 CTAs that finish early, so
        // that the others can exit and free their SMs for the next kernel, costs 9 us of combine against 3 us when all
        // 148 SMs share it: the pass is bound by per-SM L2 read latency, not by work.)
        if (lane == 0) { s.comb_rank = blockIdx.x; s.comb_total = G; }
    } else {
        // ------------------------------------------------------------------ plan (warps 2-9, into shared memory)
        if (warp < 2 + kPlanWarps) {
            const int pw = warp - 2;
            const int pt = threadIdx.x - 64;
            constexpr int kPT = kPlanWarps * 32;
            // N <= 256 (the decode sizes this kernel is tuned for) is one trip; larger batches pay one round trip each
            int ev0[I64B_MAX_TOPK];
            {
                const int t = pt;
                const long long id = (p.token_ids != nullptr && t < N) ? p.token_ids[t] : 0;
                const long long c = id < 0 ? 0 : (id >= p.vocab ? p.vocab - 1 : id);
                if (pt == 0 && (c | 1)) TL(3);                     // ids have arrived
#pragma unroll
                for (int k = 0; k < I64B_MAX_TOPK; ++k) {
                    ev0[k] = 0;
                    if (k < K && t < N) ev0[k] = p.token_ids ? (int)__ldg(p.table + (int64_t)k * p.vocab + c) : (k % E);
                }
            }
#pragma unroll
            for (int k = 0; k < I64B_MAX_TOPK; ++k)
                if (k < K && pt < N) s.expert[k][pt] = (uint8_t)min(ev0[k], E - 1);
#pragma unroll 1
            for (int t = pt + kPT; t < N; t += kPT) {
                const long long id = p.token_ids != nullptr ? p.token_ids[t] : 0;
                const long long c = id < 0 ? 0 : (id >= p.vocab ? p.vocab - 1 : id);
#pragma unroll 1
                for (int k = 0; k < K; ++k)
                    s.expert[k][t] = (uint8_t)min(p.token_ids ? (int)__ldg(p.table + (int64_t)k * p.vocab + c) : (k % E), E - 1);
            }
            for (int t = N + pt; t < ((N + 127) & ~127); t += kPT)         // padding: matches no expert
                for (int k = 0; k < K; ++k) s.expert[k][t] = 0xFF;
            if (pt == 0) TL(2);                                    // table values have arrived
            named_bar_sync(2, kPT);
            // ---- fast path: the token list (first 128 rows) and the row count of the expert of the CTA's first unit,
            // which is all its GEMM1 needs to start.  Chunks of 128 tokens in (route, token) order = merged row order;
            // warp pw owns a contiguous run of chunks; lane l holds tokens 4l..4l+3 of a chunk (one 32-bit load).
            if (own_e >= 0) {
                const uint32_t pat = (uint32_t)own_e * 0x01010101u;
                const unsigned lt_mask = (1u << lane) - 1u;
                const int nch = (N + 127) >> 7, ctot = K * nch, cpw = (ctot + kPlanWarps - 1) / kPlanWarps;
                const int c_begin = pw * cpw, c_end = min(ctot, c_begin + cpw);
                int cnt = 0;
#pragma unroll 1
                for (int c = c_begin; c < c_end; ++c) {
                    const uint32_t word = reinterpret_cast<const uint32_t*>(s.expert[c / nch])[(c % nch) * 32 + lane];
                    cnt += __popc(__vcmpeq4(word, pat)) >> 3;
                }
                cnt = warp_sum(cnt);
                if (lane == 0) s.fast_cnt[pw] = cnt;
                named_bar_sync(2, kPT);
                int running = 0, total = 0;
#pragma unroll
                for (int w = 0; w < kPlanWarps; ++w) {
                    const int cw = s.fast_cnt[w];
                    if (w < pw) running += cw;
                    total += cw;
                }
                if (pt == 0) s.rows_first = total;
#pragma unroll 1
                for (int c = c_begin; c < c_end && running < kBlockM; ++c) {
                    const uint32_t eq = __vcmpeq4(reinterpret_cast<const uint32_t*>(s.expert[c / nch])[(c % nch) * 32 + lane], pat);
                    const unsigned b0 = __ballot_sync(0xffffffffu, eq & 0xFFu), b1 = __ballot_sync(0xffffffffu, eq & 0xFF00u);
                    const unsigned b2 = __ballot_sync(0xffffffffu, eq & 0xFF0000u), b3 = __ballot_sync(0xffffffffu, eq & 0xFF000000u);
                    int pos = running + __popc(b0 & lt_mask) + __popc(b1 & lt_mask) + __popc(b2 & lt_mask) + __popc(b3 & lt_mask);
                    const int t0 = (c % nch) * 128 + lane * 4;
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if ((eq >> (8 * j)) & 1u) { if (pos < kBlockM) s.own_tok[pos] = (uint16_t)(t0 + j); ++pos; }
                    running += __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
                }
            }
        }
    }
    if (kTL && tl) {                     // bring-up: arrival times at the role barrier
        if (threadIdx.x == 0) tl[29] = gtime();
        if (threadIdx.x == 32) tl[30] = gtime();
        if (threadIdx.x == 64) tl[31] = gtime();
    }
    uint32_t tmem_base = 0;
    int n_units = 0;
    if (warp != 0) {
        tc_fence_before();
        named_bar_sync(3, kThreads);
        tc_fence_after();
        tmem_base = s.tmem_ptr;
        n_units = s.n_units;
        if (threadIdx.x == 32) { st_release_cta(&s.roles_go, 1u); TL(1); }
    }
    bool have_plan = false;

    if (warp == 0) {
        if (lane == 1 && p.next_w != nullptr && p.prefetch_mode == 1) prefetch_next_weights(p, blockIdx.x, G);
        // ------------------------------------------------------------------ weight TMA: runs ahead of everything else
        // by the depth of the ring (the weights of the next unit stream in while this one waits on a dependency)
        if (lane == 0) {
            // a ring of weights goes out at once (also under the previous kernel's tail: weights are constants) ...
            int pre_issued = 0;
            if ((int)blockIdx.x < geo.U) {
                Unit u0;
                decode_unit(p, geo, blockIdx.x, u0);
                pre_issued = u0.nkb < kStages ? u0.nkb : kStages;
                for (int kb = 0; kb < pre_issued; ++kb) { issue_b(u0, kb, kb); st_release_cta(&s.b_seq, (uint32_t)kb + 1); }
            }
            // ... then the unit table and the fast-path row count exist once the others are past the role barrier
            if (ld_acquire_cta(&s.roles_go) == 0) wait_plan_slow(&s.roles_go);
            n_units = s.n_units;
            uint32_t seq = 0;
            for (int ui = 0; ui < n_units; ++ui) {
                const Unit u = s.units[ui];
                const int rows_u = unit_rows(s, u, ui == 0, have_plan);
            const int np = passes_of(rows_u);
                for (int pass = 0; pass < np; ++pass)
                    for (int kb = 0; kb < u.nkb; ++kb, ++seq) {
                        if (seq < (uint32_t)pre_issued) continue;          // issued before the plan
                        const int stage = (int)(seq % (uint32_t)kStages);
                        mbar_wait_bounded(&s.empty_bar[stage], ((seq / (uint32_t)kStages) & 1u) ^ 1u);
                        issue_b(u, kb, stage);
                        st_release_cta(&s.b_seq, seq + 1);
                    }
            }
            if (p.next_w != nullptr && p.prefetch_mode == 2) prefetch_next_weights(p, blockIdx.x, G);
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            // c_format F32 (1<<4), A/B bf16 (1<<7, 1<<10), N>>3 at 17, M>>4 at 24
            constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kNMma >> 3) << 17) |
                                       ((uint32_t)(kBlockM >> 4) << 24);
            uint32_t seq = 0;
            int iter = 0;
            for (int ui = 0; ui < n_units; ++ui) {
                const Unit u = s.units[ui];
                const int rows_u = unit_rows(s, u, ui == 0, have_plan);
            const int np = passes_of(rows_u);
                for (int pass = 0; pass < np; ++pass, ++iter) {
                    const int as = iter & 1;
                    mbar_wait_bounded(&s.tempty_bar[as], ((iter >> 1) & 1) ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + as * kNMma;
                    for (int kb = 0; kb < u.nkb; ++kb, ++seq) {
                        const int stage = (int)(seq % (uint32_t)kStages);
                        mbar_wait_bounded(&s.full_bar[stage], (seq / (uint32_t)kStages) & 1u);
                        tc_fence_after();
                        if (kTL && tl && iter == 0 && kb < 10) tl[19 + kb] = gtime();      // bring-up: stage arrival times, first unit
                        const uint64_t a_desc = sw128_desc(sA + stage * kStageA);
                        const uint64_t b_desc = sw128_desc(sB + stage * kStageB);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        umma_commit(&s.empty_bar[stage]);
                        if (kb == u.nkb - 1) { umma_commit(&s.tfull_bar[as]); TL(iter == 0 ? 6 : 9); }
                    }
                }
            }
        }
    } else if (warp < 6) {
        // ------------------------------------------------------------------ epilogue warps
        const int q = warp & 3;
        const int r = q * 32 + lane;
        // ---- full plan (these four warps are idle until the first accumulator is complete): counts, segment starts,
        // dest_row / src_tok for every assignment.  Warp w owns experts w, w + 4, ...; lane l holds tokens 4l..4l+3 of
        // a 128-token chunk; one 4-way byte compare + four ballots rank a whole chunk in (route, token) order.
        {
            const int pw4 = warp - 2;
            const unsigned lt_mask = (1u << lane) - 1u;
            const int nch = (N + 127) >> 7;
            if (warp == 2 && lane == 0) TL(16);
#pragma unroll 1
            for (int e = pw4; e < E; e += 4) {
                const uint32_t pat = (uint32_t)e * 0x01010101u;
                int rows = 0;
#pragma unroll 1
                for (int k = 0; k < K; ++k) {
                    int c = 0;
#pragma unroll 1
                    for (int ch = 0; ch < nch; ++ch)
                        c += __popc(__vcmpeq4(reinterpret_cast<const uint32_t*>(s.expert[k])[ch * 32 + lane], pat)) >> 3;
                    c = warp_sum(c);
                    if (lane == 0) { s.route_base[k * E + e] = rows; s.offsets[k * E + e] = c; }   // offsets = counts for now
                    rows += c;
                }
                if (lane == 0) s.group_rows[e] = rows;
            }
            named_bar_sync(1, 128);
            if (warp == 2 && lane == 0) TL(17);
            const bool taps = (int)blockIdx.x == G - 1 && p.tap_routes != nullptr;   // a CTA whose epilogue warps have slack (last CTA: one late unit at most)
            // the taps live in the workspace the previous kernel of the stream may still be reading (a chain-path combine
            // or GEMM2 that has already released its dependents): nothing is stored there before it has finished
            if (taps) pdl_wait_prior();
#pragma unroll 1
            for (int e = pw4; e < E; e += 4) {
                const uint32_t pat = (uint32_t)e * 0x01010101u;
                // first merged row of expert e = rows of the experts before it (E <= 64: two values per lane)
                const int gs = warp_sum((lane < e ? s.group_rows[lane] : 0) + (lane + 32 < e ? s.group_rows[lane + 32] : 0));
                if (lane == 0) s.group_start[e] = gs;
#pragma unroll 1
                for (int k = 0; k < K; ++k) {
                    int off_k = 0;                                  // per-route exclusive offset (reference convention)
                    if (taps) {
                        off_k = warp_sum((lane < e ? s.offsets[k * E + lane] : 0) + (lane + 32 < e ? s.offsets[k * E + lane + 32] : 0));
                        if (lane == 0) { p.tap_counts[k * E + e] = s.offsets[k * E + e]; p.tap_offsets[k * E + e] = off_k; }
                    }
                    int running = gs + s.route_base[k * E + e];
#pragma unroll 1
                    for (int ch = 0; ch < nch; ++ch) {
                        const uint32_t eq = __vcmpeq4(reinterpret_cast<const uint32_t*>(s.expert[k])[ch * 32 + lane], pat);
                        const unsigned b0 = __ballot_sync(0xffffffffu, eq & 0xFFu), b1 = __ballot_sync(0xffffffffu, eq & 0xFF00u);
                        const unsigned b2 = __ballot_sync(0xffffffffu, eq & 0xFF0000u), b3 = __ballot_sync(0xffffffffu, eq & 0xFF000000u);
                        int dest = running + __popc(b0 & lt_mask) + __popc(b1 & lt_mask) + __popc(b2 & lt_mask) + __popc(b3 & lt_mask);
                        const int t0 = ch * 128 + lane * 4;
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if ((eq >> (8 * j)) & 1u) {
                                const int t = t0 + j;
                                s.dest_row[k][t] = (uint16_t)dest;
                                s.src_tok[dest] = (uint16_t)t;
                                if (taps) {
                                    p.tap_routes[(int64_t)k * N + t] = e;
                                    p.tap_dest[(int64_t)k * N + t] = dest;
                                    p.tap_scatter[(int64_t)k * N + t] = off_k + dest - gs - s.route_base[k * E + e];
                                }
                                ++dest;
                            }
                        running += __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
                    }
                }
            }
            if (pw4 == 0) {                                         // total routed rows
                const int tot = warp_sum((lane < E ? s.group_rows[lane] : 0) + (lane + 32 < E ? s.group_rows[lane + 32] : 0));
                if (lane == 0) s.group_start[E] = tot;
            }
            named_bar_sync(1, 128);
            if (warp == 2 && lane == 0) { st_release_cta(&s.plan_ready, 1u); TL(18); }
            have_plan = true;
        }
        const int R = E > 0 ? s.group_start[E] : 0;      // routed rows (= N*K)
        int iter = 0;
        for (int ui = 0; ui < n_units; ++ui) {
            const Unit u = s.units[ui];
            const int rows_u = unit_rows(s, u, ui == 0, have_plan);
            const int np = passes_of(rows_u);
            for (int pass = 0; pass < np; ++pass, ++iter) {
                const int as = iter & 1;
                mbar_wait_bounded(&s.tfull_bar[as], (iter >> 1) & 1);
                tc_fence_after();
                if (warp == 2 && lane == 0 && iter == 0) TL(14);
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * kNMma;
                const bool routed = u.type == kG1Routed || u.type == kG2Routed;
                const int row = (routed ? pass : u.m_t) * kBlockM + r;      // row inside the group
                const bool valid = row < (routed ? s.group_rows[u.g] : N);
                // Output rows leave through a per-warp staging tile: a thread holds one ROW of the accumulator, so direct
                // stores would touch 32 different lines per instruction; staged, every quarter warp writes one full
                // 128-byte line (4 lines per instruction).
                uint8_t* stg = sStg + q * (32 * kStgPitch);
                const int rows_lim = (routed ? s.group_rows[u.g] : N) - ((routed ? pass : u.m_t) * kBlockM + q * 32);   // valid rows of this warp
                const int seg = lane & 7, rsub = lane >> 3;
                (void)valid;
                if (u.type <= kG1Shared) {
                    const int n_out = routed ? p.expert_inter : p.shared_inter;
                    const int col_t = u.n_t * kBn1;                         // first intermediate column of the tile
#pragma unroll 1
                    for (int c = 0; c < kBn1 / 32; ++c) {
                        if (col_t + c * 32 >= n_out) break;
                        uint32_t gr[32], ur[32];
                        tmem_ld_32x32(taddr + c * 32, gr);
                        tmem_ld_32x32(taddr + kBn1 + c * 32, ur);
                        tmem_ld_wait();
#pragma unroll
                        for (int v4 = 0; v4 < 4; ++v4) {
                            uint32_t o[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                o[j] = pack_bf16x2(silu_f32(__uint_as_float(gr[8 * v4 + 2 * j])) * __uint_as_float(ur[8 * v4 + 2 * j]),
                                                   silu_f32(__uint_as_float(gr[8 * v4 + 2 * j + 1])) * __uint_as_float(ur[8 * v4 + 2 * j + 1]));
                            *reinterpret_cast<int4*>(stg + lane * kStgPitch + c * 64 + v4 * 16) = make_int4(o[0], o[1], o[2], o[3]);
                        }
                    }
                    __syncwarp();
                    // row pitch and first byte of the tile's columns in the intermediate buffer
                    const int64_t pitch = (int64_t)n_out * 2;
                    uint8_t* gbase = (routed ? reinterpret_cast<uint8_t*>(p.inter_r) + (int64_t)(s.group_start[u.g] + pass * kBlockM + q * 32) * pitch
                                             : reinterpret_cast<uint8_t*>(p.inter_s) + (int64_t)(u.m_t * kBlockM + q * 32) * pitch) + (int64_t)col_t * 2;
                    const bool seg_ok = col_t + seg * 8 < n_out;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int rl = rsub + 4 * i;
                        if (rl < rows_lim && seg_ok)
                            *reinterpret_cast<int4*>(gbase + rl * pitch + seg * 16) = *reinterpret_cast<const int4*>(stg + rl * kStgPitch + seg * 16);
                    }
                    __syncwarp();
                } else {
                    const int64_t yrow0 = (routed ? (int64_t)s.group_start[u.g] + pass * kBlockM
                                                  : (int64_t)R + (int64_t)u.slab * N + u.m_t * kBlockM) + q * 32;
#pragma unroll 1
                    for (int c = 0; c < kNMma / 32; ++c) {
                        const int col0 = u.n_t * kNMma + c * 32;
                        if (col0 >= H) break;
                        uint32_t ar[32];
                        tmem_ld_32x32(taddr + c * 32, ar);
                        tmem_ld_wait();
#pragma unroll
                        for (int v4 = 0; v4 < 8; ++v4)
                            *reinterpret_cast<int4*>(stg + lane * kStgPitch + v4 * 16) =
                                make_int4(ar[4 * v4], ar[4 * v4 + 1], ar[4 * v4 + 2], ar[4 * v4 + 3]);
                        __syncwarp();
                        float* gbase = p.y + yrow0 * H + col0;
                        const bool seg_ok = col0 + seg * 4 < H;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int rl = rsub + 4 * i;
                            if (rl < rows_lim && seg_ok)
                                *reinterpret_cast<int4*>(gbase + (int64_t)rl * H + seg * 4) = *reinterpret_cast<const int4*>(stg + rl * kStgPitch + seg * 16);
                        }
                        __syncwarp();
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s.tempty_bar[as]);
            }
            // publish the unit: the barrier orders the four epilogue warps' stores before one thread's gpu-scope
            // release (cumulative), so a single fence covers them all
            if (warp == 2 && lane == 0 && iter == np) TL(15);
            named_bar_sync(1, 128);
            if (warp == 2 && lane == 0) {
                if (u.type == kG1Routed) red_release_add(p.sync + I64B_SYNC_G1(u.g), 1);
                else if (u.type == kG1Shared) red_release_add(p.sync + I64B_SYNC_G1(E + u.m_t), 1);
                else red_release_add(p.sync + I64B_SYNC_G2, 1);
                TL(u.type <= kG1Shared ? 7 : 10);
            }
        }
    } else if (warp < 10) {
        // ------------------------------------------------------------------ gather warps: A rows of the routed GEMM1
        // units.  Lane l of gather warp w copies 16-byte chunk (l & 7) of tile rows 32w + 4j + (l >> 3), j = 0..7.
        const int gw = warp - 6;
        const int chunk = lane & 7;
        pdl_wait_prior();
        uint32_t seq = 0;
        const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(p.x);
        for (int ui = 0; ui < n_units; ++ui) {
            const Unit u = s.units[ui];
            const int rows_u = unit_rows(s, u, ui == 0, have_plan);
            const int np = passes_of(rows_u);
            if (u.type != kG1Routed) { seq += (uint32_t)(np * u.nkb); continue; }
            for (int pass = 0; pass < np; ++pass) {
                const __nv_bfloat16* src[8];
                uint32_t dst_off[8];
                // the CTA's first 128 rows come from the fast path (the full plan may still be in the making)
                const bool fast = ui == 0 && pass == 0;
                if (!fast && !have_plan) {
                    if (ld_acquire_cta(&s.plan_ready) == 0) wait_plan_slow(&s.plan_ready);
                    have_plan = true;
                }
                const uint16_t* toks = fast ? s.own_tok : s.src_tok + s.group_start[u.g] + pass * kBlockM;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int rj = gw * 32 + j * 4 + (lane >> 3);
                    const int tok = pass * kBlockM + rj < rows_u ? (int)toks[rj] : -1;
                    src[j] = tok >= 0 ? xb + (int64_t)tok * H + chunk * 8 : nullptr;
                    dst_off[j] = (uint32_t)(rj * 128 + ((chunk ^ (rj & 7)) << 4));
                }
                for (int kb = 0; kb < u.nkb; ++kb, ++seq) {
                    const int stage = (int)(seq % (uint32_t)kStages);
                    wait_stage_owned(&s.b_seq, seq);
                    const uint32_t dst = smem_u32(sA + stage * kStageA);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (src[j] != nullptr) cp_async_16(dst + dst_off[j], src[j] + kb * kBlockK);
                    cp_async_arrive_noinc(&s.full_bar[stage]);
                    if (kTL && tl && gw == 0 && lane == 0 && seq == 0) {          // bring-up: A-row gather latency of the first k-block
                        tl[8] = gtime();
                        asm volatile("cp.async.wait_all;" ::: "memory");
                    }
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ activation TMA (one thread):
        //   GEMM1 shared : rows of x in token order
        //   GEMM2        : rows of `inter`, after the GEMM1 units that produce them have published
        if (lane == 0) {
            pdl_wait_prior();
            uint32_t seq = 0;
            const long long t_begin = clock64();
            for (int ui = 0; ui < n_units; ++ui) {
                const Unit u = s.units[ui];
                const int rows_u = unit_rows(s, u, ui == 0, have_plan);
            const int np = passes_of(rows_u);
                if (u.type == kG1Routed) { seq += (uint32_t)(np * u.nkb); continue; }
                if (u.type >= kG2Routed) {
                    const int* ctr = p.sync + I64B_SYNC_G1(u.type == kG2Routed ? u.g : E + u.m_t);
                    const int target = u.type == kG2Routed ? geo.nt1r : geo.nt1s;
                    while (ld_acquire(ctr) < target) {
                        __nanosleep(64);                           // keep the pollers off the publishers' L2 line
                        if (clock64() - t_begin > 4000000000ll) __trap();
                    }
                    fence_proxy_async_global();
                    TL(u.type == kG2Routed ? 4 : 5);
                }
                for (int pass = 0; pass < np; ++pass)
                    for (int kb = 0; kb < u.nkb; ++kb, ++seq) {
                        const int stage = (int)(seq % (uint32_t)kStages);
                        uint8_t* dstA = sA + stage * kStageA;
                        uint64_t* bar = &s.full_bar[stage];
                        wait_stage_owned(&s.b_seq, seq);
                        mbar_expect_tx(bar, kStageA);
                        if (u.type == kG1Shared)
                            tma_load_2d(dstA, &tm_x, bar, kb * kBlockK, u.m_t * kBlockM, I64B_L2_EVICT_NORMAL);
                        else if (u.type == kG2Routed)
                            tma_load_2d(dstA, &tm_inter_r, bar, kb * kBlockK, s.group_start[u.g] + pass * kBlockM,
                                        I64B_L2_EVICT_NORMAL);
                        else
                            tma_load_2d(dstA, &tm_inter_s, bar, u.slab * kSlabK + kb * kBlockK, u.m_t * kBlockM,
                                        I64B_L2_EVICT_NORMAL);
                        mbar_arrive_n(bar, kGatherThreads);
                    }
            }
        }
    }

    // ------------------------------------------------------------------ combine (workers only)
    __syncthreads();
    const int crank = s.comb_rank, ctot = s.comb_total;
    if (crank >= 0) {
    if (threadIdx.x == 0) { TL(11); wait_counter(p.sync + I64B_SYNC_G2, geo.U2); TL(12); }
    __syncthreads();
    const int R = E > 0 ? s.group_start[E] : 0;      // routed rows (= N*K); the full plan is complete by now
    {
        // one item = four output columns of one of this worker's tokens (t = rank + j * workers); two items per trip
        // so that the loads of all of them are in flight together
        const int hv = H / 4;
        const int ntok = N > crank ? (N - crank + ctot - 1) / ctot : 0;
        const int items = ntok * hv;
        auto combine_item = [&](int item) -> float4 {
            const int t = crank + (item / hv) * ctot, c = item % hv;
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            float4 yk[I64B_MAX_TOPK];
#pragma unroll
            for (int k = 0; k < I64B_MAX_TOPK; ++k)
                if (k < K) yk[k] = __ldcg(reinterpret_cast<const float4*>(p.y + (int64_t)s.dest_row[k][t] * H) + c);
            // the slab rows are loaded four at a time before they are added (a load-add loop would serialise one L2
            // round trip per slab; a full unroll costs cold code); the sum keeps the slab order
            float4 sh = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
            for (int s0 = 0; s0 < geo.slabs; s0 += 4) {
                float4 ys[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (s0 + j < geo.slabs)
                        ys[j] = __ldcg(reinterpret_cast<const float4*>(p.y + ((int64_t)R + (int64_t)(s0 + j) * N + t) * H) + c);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (s0 + j < geo.slabs) {
                        if (s0 + j == 0) { sh = ys[0]; continue; }
                        sh.x = __fadd_rn(sh.x, ys[j].x); sh.y = __fadd_rn(sh.y, ys[j].y);
                        sh.z = __fadd_rn(sh.z, ys[j].z); sh.w = __fadd_rn(sh.w, ys[j].w);
                    }
            }
            if (E > 0) {
#pragma unroll
                for (int k = 0; k < I64B_MAX_TOPK; ++k) {
                    if (k < K) {
                        const float4 y = yk[k];
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
            if (geo.slabs > 0) {
                if (E > 0) {
                    const float ss = p.shared_scale;
                    o.x = __fadd_rn(o.x, __fmul_rn(ss, sh.x)); o.y = __fadd_rn(o.y, __fmul_rn(ss, sh.y));
                    o.z = __fadd_rn(o.z, __fmul_rn(ss, sh.z)); o.w = __fadd_rn(o.w, __fmul_rn(ss, sh.w));
                } else {
                    o = sh;
                }
            }
            return o;
        };
        auto store_item = [&](int item, const float4& o) {
            const int t = crank + (item / hv) * ctot, c = item % hv;
            if (p.out_dtype == I64B_F32) {
                reinterpret_cast<float4*>(p.out)[(int64_t)t * hv + c] = o;
            } else {
                __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
                reinterpret_cast<uint2*>(p.out)[(int64_t)t * hv + c] =
                    make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
            }
        };
#pragma unroll 1
        for (int i = threadIdx.x; i < items; i += 2 * kThreads) {
            const int i1 = i + kThreads;
            const bool two = i1 < items;
            const float4 o0 = combine_item(i);
            const float4 o1 = combine_item(two ? i1 : i);
            store_item(i, o0);
            if (two) store_item(i1, o1);
        }
    }
    }

    // ------------------------------------------------------------------ teardown: last CTA resets the counters
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) TL(13);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
    if (threadIdx.x == 0) {
        __threadfence();
        const int ticket = atomicAdd(p.sync + I64B_SYNC_FINISH, 1);
        if (ticket == G - 1) {
            for (int i = 0; i < I64B_SYNC_INTS; i += I64B_SYNC_STRIDE) p.sync[i] = 0;
            __threadfence();
        }
    }
}


// ---- bring-up timeline (env I64B_DEBUG_TIMELINE): two halves of [256 CTAs][32 slots] globaltimer stamps
constexpr int kTlInts = 256 * 32;
long long* g_tl = nullptr;
int g_tl_launch = 0;
void tl_print(const long long* d_half, int grid, long long g0_in, const char* what) {
    static long long h[kTlInts];
    cudaMemcpy(h, d_half, sizeof(h), cudaMemcpyDeviceToHost);
    long long g0 = g0_in;
    if (g0 == 0) { g0 = h[0]; for (int b = 0; b < grid; ++b) if (h[b * 32] && h[b * 32] < g0) g0 = h[b * 32]; }
    static const char* names[18] = {"start", "ids", "table", "fast", "fp0", "fpA", "fpB", "g2rA", "g2sA", "mma0",
                                    "epi0", "st0", "pub1", "mma1", "pub2", "cwait", "cgo", "end"};
    static const int order[18] = {0, 3, 2, 1, 16, 17, 18, 4, 5, 6, 14, 15, 7, 9, 10, 11, 12, 13};
    fprintf(stderr, "[i64b fused timeline: %s] ns since t0:\n  %-8s", what, "");
    for (int j = 0; j < 18; ++j) fprintf(stderr, " %6s", names[j]);
    fprintf(stderr, "\n");
    for (int b = 0; b < grid; b += 9) {
        fprintf(stderr, "  cta %3d:", b);
        for (int j = 0; j < 18; ++j) { const long long v = h[b * 32 + order[j]]; fprintf(stderr, " %6lld", v ? v - g0 : -1); }
        fprintf(stderr, "\n");
    }
    long long mx[32], mn[32];
    for (int j = 0; j < 32; ++j) { mx[j] = -(1ll << 60); mn[j] = 1ll << 60; }
    for (int b = 0; b < grid; ++b) for (int j = 0; j < 32; ++j) if (h[b * 32 + j]) { const long long v = h[b * 32 + j] - g0; if (v > mx[j]) mx[j] = v; if (v < mn[j]) mn[j] = v; }
    fprintf(stderr, "  min    :");
    for (int j = 0; j < 18; ++j) fprintf(stderr, " %6lld", mn[order[j]]);
    fprintf(stderr, "\n  max    :");
    for (int j = 0; j < 18; ++j) fprintf(stderr, " %6lld", mx[order[j]]);
    fprintf(stderr, "\n  first unit: gather kb0 issued | k-block arrival (full barrier passed) kb = 0..9 | at the role barrier: weight warp, MMA warp, plan warps:\n");
    for (int b = 0; b < grid; b += 18) {
        fprintf(stderr, "  cta %3d: %6lld |", b, h[b * 32 + 8] ? h[b * 32 + 8] - g0 : -1);
        for (int j = 19; j < 32; ++j) { const long long v = h[b * 32 + j]; fprintf(stderr, "%s %6lld", j == 29 ? " |" : "", v ? v - g0 : -1); }
        fprintf(stderr, "\n");
    }
}

constexpr int kSmemBytes = 1024 + kStages * (kStageA + kStageB) + kStgBytes + (int)sizeof(Smem) + 256;

}  // namespace

int i64b_fused_supported(int n, int hidden, int expert_inter, int shared_inter, int num_experts, int top_k) {
    if (n < 1 || n > kMaxTok) return 0;
    if (hidden % 64 != 0 || hidden > 8192) return 0;
    if (num_experts > 0 && expert_inter % 16 != 0) return 0;
    if (shared_inter % 16 != 0) return 0;
    if (num_experts + (n + kBlockM - 1) / kBlockM > I64B_FUSED_MAX_GROUPS) return 0;
    if ((int64_t)n * top_k >= 65536) return 0;             // uint16 destination rows
    if (shared_inter > kMaxSlabs * kSlabK) return 0;
    {   // the per-CTA unit table holds kMaxUnitsPerCta units (the role loops read it; larger work lists take the chain)
        I64bFusedParams q = {};
        q.n = n; q.hidden = hidden; q.expert_inter = expert_inter; q.shared_inter = shared_inter;
        q.num_experts = num_experts; q.top_k = top_k;
        if (make_geo(q).U > kMaxUnitsPerCta * i64b_num_sms()) return 0;
    }
    static_assert(kSmemBytes <= 232448, "shared memory budget");
    return 1;
}

int i64b_launch_fused(const CUtensorMap& tm_x, const CUtensorMap& tm_w13, const CUtensorMap& tm_inter_r, const CUtensorMap& tm_inter_s,
                      const CUtensorMap& tm_w2r, const CUtensorMap& tm_w2s, const I64bFusedParams& p,
                      cudaStream_t stream) {
    static bool configured[I64B_MAX_DEVICES] = {};      // function attributes are per device
    const int dev = i64b_current_device();
    if (!configured[dev]) {
        I64B_CUDA(cudaFuncSetAttribute(i64b_fused_layer_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        I64B_CUDA(cudaFuncSetAttribute(i64b_fused_layer_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        configured[dev] = true;
    }
    const int grid = i64b_num_sms();
    static const bool dbg_tl = getenv("I64B_DEBUG_TIMELINE") != nullptr;
    I64bFusedParams q = p;
    q.geo = make_geo(p);
    if (!dbg_tl) {
        I64B_CUDA(i64b_launch(i64b_fused_layer_kernel<false>, dim3(grid), dim3(kThreads), kSmemBytes, stream, tm_x, tm_w13,
                              tm_inter_r, tm_inter_s, tm_w2r, tm_w2s, q));
        i64b_count_launch(1);
        return I64B_OK;
    }
    // ---- bring-up instrumentation only (allocates + synchronises; never active unless the env asks for it)
    if (g_tl == nullptr) { cudaMalloc(&g_tl, 2 * kTlInts * sizeof(long long)); cudaMemset(g_tl, 0, 2 * kTlInts * sizeof(long long)); }
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cs);
    if (cs == cudaStreamCaptureStatusNone) {
        cudaMemsetAsync(g_tl, 0, kTlInts * sizeof(long long), stream);
        q.timeline = g_tl;
    } else {
        // inside a graph: consecutive launches alternate between the two halves, so after a replay the halves hold the
        // last two layers of the step (i64b_debug_timeline_dump prints them on one time axis)
        q.timeline = g_tl + (size_t)(g_tl_launch++ & 1) * kTlInts;
    }
    I64B_CUDA(i64b_launch(i64b_fused_layer_kernel<true>, dim3(grid), dim3(kThreads), kSmemBytes, stream, tm_x, tm_w13,
                          tm_inter_r, tm_inter_s, tm_w2r, tm_w2s, q));
    i64b_count_launch(1);
    if (cs == cudaStreamCaptureStatusNone) {
        static int printed = 0;
        cudaStreamSynchronize(stream);
        if (printed++ >= 2 && printed <= 6) tl_print(g_tl, grid, 0, "eager launch");
    }
    return I64B_OK;
}

// Bring-up only: after a graph replay under I64B_DEBUG_TIMELINE, print the last two layers of the step on one time axis.
extern "C" int i64b_debug_timeline_dump(void) {
    if (g_tl == nullptr) return 0;
    cudaDeviceSynchronize();
    static long long h[2 * kTlInts];
    cudaMemcpy(h, g_tl, sizeof(h), cudaMemcpyDeviceToHost);
    const int grid = i64b_num_sms();
    long long m[2] = {1ll << 62, 1ll << 62};
    for (int half = 0; half < 2; ++half)
        for (int b = 0; b < grid; ++b) { const long long v = h[half * kTlInts + b * 32]; if (v && v < m[half]) m[half] = v; }
    const int first = m[0] <= m[1] ? 0 : 1;
    tl_print(g_tl + (size_t)first * kTlInts, grid, m[first], "graph: second-to-last layer");
    tl_print(g_tl + (size_t)(1 - first) * kTlInts, grid, m[first], "graph: last layer (same t0)");
    return 1;
}
