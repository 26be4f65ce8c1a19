// Grouped expert GEMM for sm_100a: TMA-fed, tcgen05.mma with TMEM accumulators, persistent,
// warp-specialised (1 TMA thread, 1 MMA-issuing thread, 4 epilogue warps), device-side tile
// scheduling over the group table produced by the plan kernels (no host synchronisation).
//
//   C[rows_g, n_out_g] = A[rows_g, K] · B_g[n_out_g, K]^T          (A, B K-major; bf16→fp32 or int8→int32)
//
// Two shapes of the same kernel:
//   kDual  (GEMM1, gate|up): each stage holds a gate tile and an up tile of the packed weight, one
//          UMMA of width N = 2·BN writes both accumulators, and the epilogue fuses SiLU(gate)·up:
//            bf16 : fp32 SiLU, bf16 store           (layers/token_routed_mlp.py:148-150)
//            int8 : per-row × per-channel dequant, Q7 LUT SiLU × Q7 up → Q14 int32 store + row |max|
//                   (core/quantization.py:227-286, layers/integer_activations.py:53-66)
//   !kDual (GEMM2, down; dense linears): plain fp32/bf16 store or W8A8 dequant
//          `acc.float() * (x_scale ⊗ w_scale)` (core/quantization.py:173).
//
// Reference counterparts being replaced: cublasHgemm loop with a D2H sync per call
// (csrc/i64_expert_dispatch.cu:144-233), the 1-thread-per-output INT8 GEMMs (csrc/I64_gemm.cu:40-126),
// Triton tl.dot kernels (kernels/triton_fused_expert.py, kernels/triton/I64_int8_gemm.py) and
// torch._int_mm call sites (core/quantization.py:168-172).
#include <stdlib.h>
#include "i64b_gemm.cuh"

namespace {

constexpr int kBlockM = 128;
constexpr int kRowBytes = 128;                     // one 128B swizzle span of K per stage row
constexpr int kStageA = kBlockM * kRowBytes;       // 16 KiB
constexpr int kThreads = 320;                      // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue
constexpr int kEpiWarps = 8;
constexpr int kMiscBytes = 8192;
constexpr int kStagePitch = 80;                    // 64-byte row payload + 16: row-per-lane 16 B stores hit 8 distinct bank groups
constexpr int kStageWarpBytes = 32 * kStagePitch;  // per-warp output staging block (32 rows x 64 B)
constexpr int kScaleWarpBytes = 512;               // per-warp column scales of the current tile (<= 128 floats)
constexpr int kEpiWarpBytes = kStageWarpBytes + kScaleWarpBytes;

template <int kNMma, int kCtas> struct StageCfg {
    static constexpr int kStageB = kNMma / kCtas * kRowBytes;     // a CTA pair splits the B tile
    static constexpr int kStagesRaw = 196608 / (kStageA + kStageB);
    static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
    static constexpr int kSmem = 1024 + kStages * (kStageA + kStageB) + kMiscBytes + kEpiWarps * kEpiWarpBytes;
    static_assert(kSmem <= 232448, "shared memory budget (227 KB per CTA)");
};

__device__ __forceinline__ uint64_t make_sw128_kmajor_desc(const void* smem_ptr) {
    uint64_t d = (uint64_t)((smem_u32(smem_ptr) >> 4) & 0x3FFFu);   // start address
    d |= (uint64_t)(1024 >> 4) << 32;                                // SBO: 8 rows x 128 B
    d |= 1ull << 46;                                                 // descriptor version (sm_100)
    d |= 2ull << 61;                                                 // SWIZZLE_128B
    return d;
}

// operand kinds of the grouped GEMM
constexpr int kTypeBf16 = I64B_TYPE_BF16, kTypeInt8 = I64B_TYPE_INT8, kTypeFp8 = I64B_TYPE_FP8;

template <int kType, int kNMma, int kCtas>
__device__ __forceinline__ constexpr uint32_t make_idesc() {
    // c_format [4,6): 1=F32, 2=S32; a/b format [7,10)/[10,13): kind::f16 BF16=1 / kind::i8 S8=1 / kind::f8f6f4 E4M3=0;
    // K-major both; n_dim [17,23) = N>>3; m_dim [24,29) = M>>4
    constexpr uint32_t ab = kType == kTypeFp8 ? 0u : 1u;
    return (uint32_t)((kType == kTypeInt8 ? 2u : 1u) << 4) | (ab << 7) | (ab << 10) | ((uint32_t)(kNMma >> 3) << 17) |
           ((uint32_t)((kBlockM * kCtas) >> 4) << 24);
}

struct TileCoord {
    int g, m_t, n_t;
};

// `tile` counts (m-tile x kCtas, n-tile) units; CTA `rank` of a pair takes the rank-th 128-row tile of the unit
template <int kBnOut, int kCtas>
__device__ __forceinline__ void decode_tile(int tile, int& g, const int* prefix, const I64bGroup* sg,
                                            TileCoord& tc, int rank) {
    while (tile >= prefix[g + 1]) ++g;
    const int nt = (sg[g].n_out + kBnOut - 1) / kBnOut;
    const int local = tile - prefix[g];
    tc.g = g;
    tc.m_t = (local / nt) * kCtas + rank;
    tc.n_t = local % nt;
}

__device__ __forceinline__ float silu_f32(float g) {
    // branch-free: ex2.approx + rcp.approx (an IEEE rcp carries a slow-path branch per element, which serialises
    // the 32-element chunk on the single epilogue warp of each SM sub-partition)
    return __fdividef(g, 1.0f + __expf(-g));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void st_f32(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" :: "r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ int ld_shared_s16(uint32_t addr) {
    int v;
    asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
// One [32 rows x 64 bytes] output block of a warp: lane i supplies the 64 bytes of row i (the TMEM read layout);
// the block is transposed through the warp's staging area and leaves as 64-byte row segments, 8 rows per store
// instruction (16 full sectors) instead of 32 rows x 16 bytes.  Rows >= rows_valid and bytes >= bytes_valid
// (a multiple of 16) are not written.
__device__ __forceinline__ void warp_store_block(uint32_t st_addr, const uint32_t (&v)[16], uint8_t* gbase,
                                                 int64_t pitch_bytes, int rows_valid, int bytes_valid, int lane) {
    __syncwarp();                                   // the previous block's reads are done
#pragma unroll
    for (int j = 0; j < 4; ++j)
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};"
                     :: "r"(st_addr + lane * kStagePitch + j * 16), "r"(v[4 * j]), "r"(v[4 * j + 1]), "r"(v[4 * j + 2]),
                        "r"(v[4 * j + 3]) : "memory");
    __syncwarp();
    const int piece = lane & 3;
    const int rr = (lane >> 3) + 4 * ((lane >> 2) & 1);   // quarter-warps read rows {k, k+4}: 16 banks apart
    int4 d[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)         // all four reads first: one shared-memory latency per block, not four
        asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(d[i].x), "=r"(d[i].y), "=r"(d[i].z), "=r"(d[i].w)
                     : "r"(st_addr + (i * 8 + rr) * kStagePitch + piece * 16));
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = i * 8 + rr;
        if (row < rows_valid && piece * 16 < bytes_valid)
            *reinterpret_cast<int4*>(gbase + row * pitch_bytes + piece * 16) = d[i];
    }
}

// The same block with every row at its own address: lane i supplies `my_off`, the byte offset of row i from gbase.
__device__ __forceinline__ void warp_store_block_rows(uint32_t st_addr, const uint32_t (&v)[16], uint8_t* gbase,
                                                      int64_t my_off, int rows_valid, int bytes_valid, int lane) {
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 4; ++j)
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};"
                     :: "r"(st_addr + lane * kStagePitch + j * 16), "r"(v[4 * j]), "r"(v[4 * j + 1]), "r"(v[4 * j + 2]),
                        "r"(v[4 * j + 3]) : "memory");
    __syncwarp();
    const int piece = lane & 3;
    const int rr = (lane >> 3) + 4 * ((lane >> 2) & 1);
    int4 d[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(d[i].x), "=r"(d[i].y), "=r"(d[i].z), "=r"(d[i].w)
                     : "r"(st_addr + (i * 8 + rr) * kStagePitch + piece * 16));
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = i * 8 + rr;
        const int64_t off = __shfl_sync(0xffffffffu, my_off, row);
        if (row < rows_valid && piece * 16 < bytes_valid)
            *reinterpret_cast<int4*>(gbase + off + piece * 16) = d[i];
    }
}

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_cg_f4(const float* p) {      // L2-coherent load (rows written by other CTAs)
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

// Fused combine for one [32 tokens x 16 columns] block of a shared-expert GEMM2 tile.  Lane i supplies the shared
// expert's fp32 outputs of token row i; after the transpose through the staging area each lane owns 4 consecutive
// columns of 4 token rows, adds the K routed rows of those tokens (loaded by the caller, a whole 32-column chunk in
// flight before the accumulator is read) and writes the finished block output in the reference's operation order
// (no FMA contraction):   o = w0*y0; o = o + wk*yk; o = rs*o; o = o + ss*shared     layers/token_routed_mlp.py:156-185
constexpr int kFuseMaxK = 2;     // the fused combine covers top_k <= 2 (the named models); wider routing takes combine_kernel
struct CombineCtx {
    const float* y;          // routed per-assignment rows
    int64_t ldy;
    int top_k;
    float w[kFuseMaxK];
    float rs, ss;
    int prk[4][kFuseMaxK];   // routed row of (token row i*8+rr, route k)
};
// issue the loads of the routed rows for one 16-column block (4 token rows x K routes per lane)
__device__ __forceinline__ void combine_issue_loads(const CombineCtx& cx, int col, int rows_valid, int lane,
                                                    float4 (&yk)[4][kFuseMaxK]) {
    const int piece = lane & 3;
    const int rr = (lane >> 3) + 4 * ((lane >> 2) & 1);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const bool rv = i * 8 + rr < rows_valid;
#pragma unroll
        for (int k = 0; k < kFuseMaxK; ++k)
            yk[i][k] = (rv && k < cx.top_k) ? ld_cg_f4(cx.y + (int64_t)cx.prk[i][k] * cx.ldy + col + piece * 4)
                                           : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
__device__ __forceinline__ void warp_combine_block(uint32_t st_addr, const uint32_t (&v)[16], const CombineCtx& cx,
                                                   const float4 (&yk)[4][kFuseMaxK], uint8_t* obase, int64_t opitch,
                                                   bool f32out, uint8_t* ykeep, int64_t ykeep_pitch, int rows_valid,
                                                   int lane) {
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 4; ++j)
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};"
                     :: "r"(st_addr + lane * kStagePitch + j * 16), "r"(v[4 * j]), "r"(v[4 * j + 1]), "r"(v[4 * j + 2]),
                        "r"(v[4 * j + 3]) : "memory");
    __syncwarp();
    const int piece = lane & 3;
    const int rr = (lane >> 3) + 4 * ((lane >> 2) & 1);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = i * 8 + rr;
        float4 a;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(st_addr + row * kStagePitch + piece * 16));
        float4 o = a;                    // dense FFN (no routed experts): the SwiGLU output itself
        if (cx.top_k > 0) {
            const float4 y0 = yk[i][0];
            o = make_float4(__fmul_rn(cx.w[0], y0.x), __fmul_rn(cx.w[0], y0.y), __fmul_rn(cx.w[0], y0.z),
                            __fmul_rn(cx.w[0], y0.w));
            if (cx.top_k > 1) {
                const float4 y1 = yk[i][1];
                o.x = __fadd_rn(o.x, __fmul_rn(cx.w[1], y1.x));
                o.y = __fadd_rn(o.y, __fmul_rn(cx.w[1], y1.y));
                o.z = __fadd_rn(o.z, __fmul_rn(cx.w[1], y1.z));
                o.w = __fadd_rn(o.w, __fmul_rn(cx.w[1], y1.w));
            }
            o = make_float4(__fmul_rn(cx.rs, o.x), __fmul_rn(cx.rs, o.y), __fmul_rn(cx.rs, o.z), __fmul_rn(cx.rs, o.w));
            o.x = __fadd_rn(o.x, __fmul_rn(cx.ss, a.x));
            o.y = __fadd_rn(o.y, __fmul_rn(cx.ss, a.y));
            o.z = __fadd_rn(o.z, __fmul_rn(cx.ss, a.z));
            o.w = __fadd_rn(o.w, __fmul_rn(cx.ss, a.w));
        }
        if (row < rows_valid) {
            if (f32out) {
                *reinterpret_cast<float4*>(obase + row * opitch + piece * 16) = o;
            } else {
                *reinterpret_cast<uint2*>(obase + row * opitch + piece * 8) =
                    make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
            }
            if (ykeep != nullptr) *reinterpret_cast<float4*>(ykeep + row * ykeep_pitch + piece * 16) = a;
        }
    }
}

template <int kType, bool kDual, int kNMma, int kCtas, bool kGather = false>
__global__ void __launch_bounds__(kThreads, 1)
i64b_gemm_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                 const __grid_constant__ CUtensorMap tmB0, const __grid_constant__ CUtensorMap tmB1,
                 const __grid_constant__ I64bGemmParams p) {
    constexpr bool kInt8 = kType == kTypeInt8;
    constexpr bool kFp8 = kType == kTypeFp8;      // E4M3 operands, fp32 accumulators, per-row x per-channel scales
    using Cfg = StageCfg<kNMma, kCtas>;
    static_assert(kCtas == 1 || kCtas == 2, "single CTA or a CTA pair");
    static_assert(kCtas == 1 || kNMma == 256, "the pair kernel is built for the widest tile only");
    constexpr int kStages = Cfg::kStages;
    constexpr int kStageB = Cfg::kStageB;
    constexpr int kBnOut = kDual ? kNMma / 2 : kNMma;
    constexpr int kBlockK = kType != kTypeBf16 ? 128 : 64;   // elements per 128-byte row
    constexpr uint32_t kTmemCols = 2 * kNMma;          // two accumulator stages (power of two ≥ 32)
    static_assert((kDual ? kNMma / 2 : kNMma) >= 32, "a warp drains at least one 32-column TMEM chunk");
    static_assert(kTmemCols >= 32 && kTmemCols <= 512 && (kTmemCols & (kTmemCols - 1)) == 0, "TMEM columns");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + kStages * kStageA;
    uint8_t* misc = sB + kStages * kStageB;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(misc);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tfull_bar = empty_bar + kStages;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    int* prefix = reinterpret_cast<int*>(tmem_ptr + 2);
    I64bGroup* sg = reinterpret_cast<I64bGroup*>(misc + 512);
    int16_t* slut = reinterpret_cast<int16_t*>(misc + 512 + sizeof(I64bGroup) * I64B_MAX_GROUPS);
    static_assert(512 + sizeof(I64bGroup) * I64B_MAX_GROUPS + 2 * I64B_SILU_LUT_SIZE <= kMiscBytes, "misc smem");
    uint8_t* stage_base = misc + kMiscBytes;
    static_assert((2 * 8 + 2 * 2) * 8 + 8 + 4 * (I64B_MAX_GROUPS + 1) <= 512, "barrier area");

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int rank = kCtas == 2 ? (int)cluster_ctarank() : 0;     // 0 = leader of the pair (issues the MMAs)
    const int unit0 = (int)blockIdx.x / kCtas;                    // first tile unit of this CTA (pair)
    const int unit_step = (int)gridDim.x / kCtas;
    pdl_launch_dependents();      // let the next kernel's CTAs start their prologue as SMs free up
    const long long t0 = clock64();
    long long* tl = p.debug_timeline ? p.debug_timeline + (size_t)blockIdx.x * 8 : nullptr;
    if (tl && threadIdx.x == 0) {
        unsigned long long gt;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        tl[0] = (long long)gt;
    }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmB0);
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmB1);
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], kEpiWarps * kCtas);     // the leader's copy collects both CTAs' epilogue warps
        }
        fence_barrier_init();
    } else if (warp == 1) {
        if constexpr (kCtas == 2) tmem_alloc_cta2<kTmemCols>(tmem_ptr);
        else tmem_alloc<kTmemCols>(tmem_ptr);
    }
    if constexpr (kCtas == 2) cluster_sync_all();     // the peer's barriers exist before anything signals them

    // everything above overlaps the previous kernel's tail under PDL; below reads its outputs
    pdl_wait_prior();

    const int ng = p.num_groups;
    for (int i = threadIdx.x; i < ng; i += kThreads) sg[i] = p.groups ? p.groups[i] : p.inline_group;
    if constexpr (kInt8 && kDual) {
        for (int i = threadIdx.x; i < I64B_SILU_LUT_SIZE; i += kThreads) slut[i] = p.silu_lut[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int g = 0; g < ng; ++g) {
            prefix[g] = acc;
            const int mt = (sg[g].rows + kBlockM * kCtas - 1) / (kBlockM * kCtas);
            const int nt = (sg[g].n_out + kBnOut - 1) / kBnOut;
            acc += mt * nt;
        }
        prefix[ng] = acc;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const int total_tiles = prefix[ng];
    if (tl && threadIdx.x == 0) tl[1] = clock64() - t0;

    if (warp == 0 && kGather) {
        // ------------------------------------------------------------------ TMA producer, routed A rows gathered from x
        // (no dispatched copy): the whole warp runs the loop; per k-block lane l fetches rows 4l..4l+3 of the tile with one
        // tile::gather4 (4 x 128 bytes, landing where a box load would have put them), lane 0 adds the weight half-tile.
        static_assert(!kGather || (kCtas == 2 && kDual && !kInt8), "gather: bf16 GEMM1 pair kernel");
        int stage = 0, g = 0;
        uint32_t phase = 0;
        for (int tile = unit0; tile < total_tiles; tile += unit_step) {
            TileCoord tc;
            decode_tile<kBnOut, kCtas>(tile, g, prefix, sg, tc, rank);
            const I64bGroup grp = sg[tc.g];
            const bool gat = grp.which == 0;
            const int nkb = (p.k[grp.which] + kBlockK - 1) / kBlockK;
            const int a_row = grp.row_start + tc.m_t * kBlockM;
            const int b_row = grp.w_row_base + tc.n_t * kBnOut;
            int ix[4] = {0, 0, 0, 0};
            if (gat) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = a_row + 4 * lane + j;
                    const int t = r < p.gather_rows ? __ldg(p.gather_idx + r) : 0;
                    ix[j] = min(max(t, 0), p.n_tokens - 1);
                }
            }
            for (int kb = 0; kb < nkb; ++kb) {
                if (lane == 0) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * (kStageA + kStageB));
                }
                __syncwarp();
                if (gat) {
                    tma_gather4_cta2(sA + stage * kStageA + lane * 512, &tmA0, &full_bar[stage], kb * kBlockK, ix[0], ix[1],
                                     ix[2], ix[3], I64B_L2_EVICT_NORMAL);
                } else if (lane == 0) {
                    tma_load_2d_cta2(sA + stage * kStageA, &tmA1, &full_bar[stage], kb * kBlockK, a_row, I64B_L2_EVICT_NORMAL);
                }
                if (lane == 0) {
                    const int bh = b_row + rank * grp.n_out;
                    tma_load_2d_cta2(sB + stage * kStageB, grp.which ? &tmB1 : &tmB0, &full_bar[stage], kb * kBlockK, bh,
                                     I64B_L2_EVICT_NORMAL);
                }
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0, g = 0;
            uint32_t phase = 0;
            for (int tile = unit0; tile < total_tiles; tile += unit_step) {
                TileCoord tc;
                decode_tile<kBnOut, kCtas>(tile, g, prefix, sg, tc, rank);
                const I64bGroup grp = sg[tc.g];
                const CUtensorMap* ta = grp.which ? &tmA1 : &tmA0;
                const CUtensorMap* tb = grp.which ? &tmB1 : &tmB0;
                const int nkb = (p.k[grp.which] + kBlockK - 1) / kBlockK;
                const int a_row = grp.row_start + tc.m_t * kBlockM;
                const int b_row = grp.w_row_base + tc.n_t * kBnOut;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (tl && kb == 0 && tile == unit0) tl[2] = clock64() - t0;
                    if constexpr (kCtas == 2) {
                        // both CTAs load their own A rows and their half of the B tile (dual: leader = gate rows,
                        // peer = up rows); all bytes are credited to the leader's barrier
                        // bring-up (timing only, results are garbage): bit7 / bit8 = no B / no A load on odd k-blocks, i.e.
                        // what the kernel would do with three quarters of its operand bytes (is it bound by operand ingest?)
                        const bool skip_b = (p.debug_flags & 128u) && (kb & 1), skip_a = (p.debug_flags & 256u) && (kb & 1);
                        if (rank == 0)
                            mbar_arrive_expect_tx(&full_bar[stage], 2 * ((skip_a ? 0 : kStageA) + (skip_b ? 0 : kStageB)));
                        if (!skip_a)
                            tma_load_2d_cta2(sA + stage * kStageA, ta, &full_bar[stage], kb * kBlockK, a_row,
                                             I64B_L2_EVICT_NORMAL);
                        const int bh = kDual ? b_row + rank * grp.n_out : b_row + rank * (kNMma / 2);
                        if (!skip_b)
                            tma_load_2d_cta2(sB + stage * kStageB, tb, &full_bar[stage], kb * kBlockK, bh,
                                             I64B_L2_EVICT_NORMAL);
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    mbar_arrive_expect_tx(&full_bar[stage], kStageA + kStageB);
                    tma_load_2d(sA + stage * kStageA, ta, &full_bar[stage], kb * kBlockK, a_row,
                                I64B_L2_EVICT_NORMAL);
                    if constexpr (kDual) {
                        tma_load_2d(sB + stage * kStageB, tb, &full_bar[stage], kb * kBlockK, b_row,
                                    I64B_L2_EVICT_NORMAL);
                        tma_load_2d(sB + stage * kStageB + kBnOut * kRowBytes, tb, &full_bar[stage],
                                    kb * kBlockK, b_row + grp.n_out, I64B_L2_EVICT_NORMAL);
                    } else {
                        tma_load_2d(sB + stage * kStageB, tb, &full_bar[stage], kb * kBlockK, b_row,
                                    I64B_L2_EVICT_NORMAL);
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (one thread)
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = make_idesc<kType, kNMma, kCtas>();
            int stage = 0, g = 0, iter = 0;
            uint32_t phase = 0;
            for (int tile = unit0; tile < total_tiles; tile += unit_step, ++iter) {
                TileCoord tc;
                decode_tile<kBnOut, kCtas>(tile, g, prefix, sg, tc, rank);
                const int nkb = (p.k[sg[tc.g].which] + kBlockK - 1) / kBlockK;
                const int as = iter & 1;
                mbar_wait(&tempty_bar[as], ((iter >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * kNMma;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    if (tl && kb == 0 && tile == unit0) tl[3] = clock64() - t0;
                    if ((p.debug_flags & 1u) && kCtas == 1) {      // pure TMA streaming: release the slot without touching the tensor core
                        mbar_arrive(&empty_bar[stage]);
                        if (kb == nkb - 1) { mbar_arrive(&tfull_bar[as]); if (tl) tl[4] = clock64() - t0; }
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    const uint64_t a_desc = make_sw128_kmajor_desc(sA + stage * kStageA);
                    const uint64_t b_desc = make_sw128_kmajor_desc(sB + stage * kStageB);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {   // 4 x 32-byte K slices inside the swizzle span
                        if constexpr (kCtas == 2) {
                            if constexpr (kInt8)
                                umma_i8_cta2(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                            else if constexpr (kFp8)
                                umma_f8_cta2(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                            else
                                umma_bf16_cta2(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        } else if constexpr (kInt8) {
                            umma_i8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        } else if constexpr (kFp8) {
                            umma_f8(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        } else {
                            umma_bf16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        }
                    }
                    if constexpr (kCtas == 2) {     // the commits arrive on both CTAs' barriers
                        umma_commit_cta2(&empty_bar[stage]);
                        if (kb == nkb - 1) umma_commit_cta2(&tfull_bar[as]);
                    } else {
                        umma_commit(&empty_bar[stage]);             // smem slot reusable once these MMAs retire
                        if (kb == nkb - 1) { umma_commit(&tfull_bar[as]); if (tl) tl[4] = clock64() - t0; }   // accumulator complete
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue (8 warps: 4 TMEM lane quarters x 2 column halves)
        // Every output block goes registers -> per-warp smem staging (row per lane, conflict-free 80-byte pitch) ->
        // coalesced 64-byte row segments in global memory (8 rows per warp instruction) instead of one row per lane.
        const int q = warp & 3;                 // TMEM lane quarter this warp may access (hardware: warp id % 4)
        const int h = (warp - 2) >> 2;          // column half of the tile this warp drains
        const int r = q * 32 + lane;            // tile row owned by this thread while reading TMEM
        constexpr bool kSplit = kBnOut >= 64;   // narrow tiles: only the h == 0 warps work
        constexpr int kColsW = kSplit ? kBnOut / 2 : kBnOut;    // output columns per warp
        const bool active = kSplit || h == 0;
        const int cbase = kSplit ? h * kColsW : 0;
        const uint32_t st_addr = smem_u32(stage_base + (warp - 2) * kEpiWarpBytes);
        float* wscale = reinterpret_cast<float*>(stage_base + (warp - 2) * kEpiWarpBytes + kStageWarpBytes);
        const uint32_t ws_addr = st_addr + kStageWarpBytes;
        const uint32_t lut_addr = smem_u32(slut);
        int g = 0, iter = 0;
        [[maybe_unused]] bool routed_ready = false;
        for (int tile = unit0; tile < total_tiles; tile += unit_step, ++iter) {
            TileCoord tc;
            decode_tile<kBnOut, kCtas>(tile, g, prefix, sg, tc, rank);
            const I64bGroup grp = sg[tc.g];
            const int w = grp.which;
            const int as = iter & 1;
            const int n_out = grp.n_out;
            const int row_in_group = tc.m_t * kBlockM + r;
            const bool valid = row_in_group < grp.rows;
            const int64_t a_row = (int64_t)grp.row_start + row_in_group;
            int64_t out_row0 = (int64_t)grp.out_row_start + tc.m_t * kBlockM + q * 32;   // first row of this warp
            if (p.debug_flags & 32u) out_row0 &= 127;       // bring-up: bit5 = all stores land in the first 128 rows (L2-resident)
            const int rows_valid = min(32, max(0, grp.rows - (tc.m_t * kBlockM + q * 32)));
            const int colw0 = tc.n_t * kBnOut + cbase;      // first output column of this warp

            // per-column scales of this warp's slice -> smem (before waiting for the accumulator)
            if constexpr (kInt8 || kFp8 || !kDual) {
                const bool need = (kInt8 || kFp8) ? (kDual || p.epi == I64B_EPI_DEQUANT) : (p.epi == I64B_EPI_DEQUANT);
                if (active && need) {
                    const float* cs = p.col_scale[w] + grp.w_row_base;
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < (kColsW + 31) / 32; ++i) {
                        const int c = i * 32 + lane;
                        if (c < kColsW) {
                            const bool cv = colw0 + c < n_out;
                            wscale[c] = cv ? __ldg(cs + colw0 + c) : 0.0f;
                            if constexpr (kDual) wscale[kColsW + c] = cv ? __ldg(cs + n_out + colw0 + c) : 0.0f;
                        }
                    }
                    __syncwarp();
                }
            }

            [[maybe_unused]] CombineCtx cx;
            bool combine_tile = false;
            if constexpr (!kDual) {
                combine_tile = p.fuse_combine != 0 && w == 1 && active;
                if (combine_tile) {
                    cx.y = reinterpret_cast<const float*>(p.out[0]);
                    cx.ldy = p.ld_out[0];
                    cx.top_k = p.top_k;
#pragma unroll
                    for (int k = 0; k < kFuseMaxK; ++k) cx.w[k] = p.w[k];
                    cx.rs = p.routed_scale;
                    cx.ss = p.shared_scale;
                    if (p.top_k > 0) {
                        // every routed tile (all groups before the shared one) must have left its epilogue
                        if (!routed_ready) {
                            if (lane == 0) {
                                const int target = prefix[ng - 1] * kEpiWarps * kCtas;
                                unsigned spins = 0;
                                while (ld_acquire_gpu(p.tile_done) < target) {
                                    __nanosleep(64);
                                    if (++spins > (1u << 26)) __trap();
                                }
                            }
                            __syncwarp();
                        }
                        const int tok = row_in_group;           // shared group rows are tokens, in order
                        int rk[kFuseMaxK];
#pragma unroll
                        for (int k = 0; k < kFuseMaxK; ++k) {
                            rk[k] = (k < p.top_k && valid) ? __ldg(p.dest_row + (int64_t)k * p.n_tokens + tok) : 0;
                            if (!routed_ready && k < p.top_k && valid) {      // first shared tile: nobody prefetched for it
                                const float* seg = cx.y + (int64_t)rk[k] * cx.ldy + colw0;
#pragma unroll
                                for (int j = 0; j < kColsW / 32; ++j) prefetch_l2(seg + j * 32);
                            }
                        }
                        routed_ready = true;
                        const int rr = (lane >> 3) + 4 * ((lane >> 2) & 1);
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int k = 0; k < kFuseMaxK; ++k)
                                cx.prk[i][k] = __shfl_sync(0xffffffffu, rk[k], i * 8 + rr);
                        // pull the routed rows of this CTA's NEXT tile into L2 now: a whole tile of MMAs ahead of their use
                        const int nxt = tile + unit_step;
                        if (nxt < total_tiles && !(p.debug_flags & 16u)) {                  // bring-up: bit4 = no prefetch-ahead
                            int g2 = g;
                            TileCoord tn;
                            decode_tile<kBnOut, kCtas>(nxt, g2, prefix, sg, tn, rank);
                            const int tok2 = tn.m_t * kBlockM + r;
                            if (sg[tn.g].which == 1 && tok2 < sg[tn.g].rows) {
#pragma unroll
                                for (int k = 0; k < kFuseMaxK; ++k) {
                                    if (k < p.top_k) {
                                        const int row2 = __ldg(p.dest_row + (int64_t)k * p.n_tokens + tok2);
                                        const float* seg = cx.y + (int64_t)row2 * cx.ldy + tn.n_t * kBnOut + cbase;
#pragma unroll
                                        for (int j = 0; j < kColsW / 32; ++j) prefetch_l2(seg + j * 32);
                                    }
                                }
                            }
                        }
                    }
                }
            }

            mbar_wait(&tfull_bar[as], (iter >> 1) & 1);
            tc_fence_after();
            if (tl && threadIdx.x == 64) tl[5] = clock64() - t0;
            if ((p.debug_flags & 2u) || !active) {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (kCtas == 2) mbar_arrive_leader(&tempty_bar[as]); else mbar_arrive(&tempty_bar[as]);
                    if constexpr (!kDual)
                        if (p.fuse_combine && w == 0) red_release_gpu_add(p.tile_done, 1);   // keep the count uniform
                }
                continue;
            }
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * kNMma + cbase;

            if constexpr (kDual && !kInt8) {
                uint8_t* obase = reinterpret_cast<uint8_t*>(p.out[w]) + (out_row0 * p.ld_out[w] + colw0) * 2;
                const int64_t pitch = p.ld_out[w] * 2;
                // token-major layout: this lane's row goes to its token's row of the [N, 2I+S] buffer
                const bool scatter = p.rowmap[w] != nullptr;
                int64_t my_off = 0;
                [[maybe_unused]] float rs8 = 0.0f;          // FP8: this row's activation scale
                if constexpr (kFp8) rs8 = valid ? p.row_scale[w][a_row] : 0.0f;
                float mul = 1.0f;           // one-accumulator down projection: w_k * rs (or ss) folded in here, a power of two
                const bool use_mul = scatter && p.tri_scaled != 0;
                if (scatter) {
                    obase = reinterpret_cast<uint8_t*>(p.out[w]);
                    if (valid) {
                        // routed rows: bit 30 = K segment of the pair-order row (0: the pair's lower expert), bit 31 = route
                        const uint32_t m = (uint32_t)__ldg(p.rowmap[w] + out_row0 + lane);
                        const int col = w ? p.tri_col[2] : p.tri_col[(m >> 30) & 1u];
                        mul = w ? p.tri_mul[2] : p.tri_mul[m >> 31];
                        my_off = ((int64_t)(m & 0x3FFFFFFFu) * p.ld_out[w] + col + colw0) * 2;
                    }
                }
#pragma unroll 1
                for (int c = 0; c < kColsW; c += 32) {
                    if (colw0 + c >= n_out) break;
                    uint32_t o[16];
                    {
                        uint32_t gr[32], ur[32];
                        tmem_ld_32x32(taddr + c, gr);
                        tmem_ld_32x32(taddr + kBnOut + c, ur);
                        tmem_ld_wait();
                        if constexpr (kFp8) {       // _scaled_mm: acc * (x_scale * w_scale), gate and up         core/fp8.py:231-245
#pragma unroll
                            for (int j4 = 0; j4 < 8; ++j4) {
                                const float4 sg4 = ld_shared_f4(ws_addr + (c + j4 * 4) * 4);
                                const float4 su4 = ld_shared_f4(ws_addr + (kColsW + c + j4 * 4) * 4);
                                const float sgv[4] = {sg4.x, sg4.y, sg4.z, sg4.w};
                                const float suv[4] = {su4.x, su4.y, su4.z, su4.w};
#pragma unroll
                                for (int jj = 0; jj < 4; ++jj) {
                                    const int j = j4 * 4 + jj;
                                    gr[j] = __float_as_uint(__fmul_rn(__uint_as_float(gr[j]), __fmul_rn(rs8, sgv[jj])));
                                    ur[j] = __float_as_uint(__fmul_rn(__uint_as_float(ur[j]), __fmul_rn(rs8, suv[jj])));
                                }
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            float h0 = silu_f32(__uint_as_float(gr[2 * j])) * __uint_as_float(ur[2 * j]);
                            float h1 = silu_f32(__uint_as_float(gr[2 * j + 1])) * __uint_as_float(ur[2 * j + 1]);
                            if (use_mul) {
                                h0 = __fmul_rn(h0, mul);
                                h1 = __fmul_rn(h1, mul);
                            }
                            o[j] = pack_bf16x2(h0, h1);
                        }
                    }
                    if (scatter)
                        warp_store_block_rows(st_addr, o, obase + c * 2, my_off, rows_valid, min(64, (n_out - colw0 - c) * 2), lane);
                    else
                        warp_store_block(st_addr, o, obase + c * 2, pitch, rows_valid, min(64, (n_out - colw0 - c) * 2), lane);
                }
            } else if constexpr (kDual && kInt8) {
                uint8_t* obase = reinterpret_cast<uint8_t*>(p.out[w]) + (out_row0 * p.ld_out[w] + colw0) * 4;
                const int64_t pitch = p.ld_out[w] * 4;
                // fold the Q7 factor into the row scale: fl(fl(a*fl(rs*s))*128) == fl(a*fl((128*rs)*s)) exactly
                // (scaling by 2^7 commutes with round-to-nearest; scales are >= 1e-8/127, nothing denormal)
                const float rs128 = valid ? __fmul_rn(p.row_scale[w][a_row], 128.0f) : 0.0f;
                int amax = 0;
#pragma unroll 1
                for (int c = 0; c < kColsW; c += 32) {
                    if (colw0 + c >= n_out) break;
                    {
                        uint32_t gr[32], ur[32];
                        tmem_ld_32x32(taddr + c, gr);
                        tmem_ld_32x32(taddr + kBnOut + c, ur);
                        tmem_ld_wait();
                        int o[32];
#pragma unroll
                        for (int j4 = 0; j4 < 8; ++j4) {
                            const float4 sg4 = ld_shared_f4(ws_addr + (c + j4 * 4) * 4);
                            const float4 su4 = ld_shared_f4(ws_addr + (kColsW + c + j4 * 4) * 4);
                            const float sgv[4] = {sg4.x, sg4.y, sg4.z, sg4.w};
                            const float suv[4] = {su4.x, su4.y, su4.z, su4.w};
#pragma unroll
                            for (int jj = 0; jj < 4; ++jj) {
                                const int j = j4 * 4 + jj;
                                // result_i32.float() * (x_scale * w_scale), then (.. * 128).round()
                                //                          core/quantization.py:278, layers/integer_activations.py:60-63
                                const int gq = __float2int_rn(__fmul_rn((float)(int)gr[j], __fmul_rn(rs128, sgv[jj])));
                                const int uq = __float2int_rn(__fmul_rn((float)(int)ur[j], __fmul_rn(rs128, suv[jj])));
                                // silu_integer: LUT on [-1024, 1024], identity above, 0 below (LUT ends are 0 and 1024,
                                // checked at load time), so the two selects collapse into one add
                                const int idx = min(max(gq, -1024), 1024);
                                const int sq = ld_shared_s16(lut_addr + (uint32_t)(idx + 1024) * 2) + max(gq - 1024, 0);
                                o[j] = sq * uq;
                            }
                        }
                        const bool second = colw0 + c + 16 < n_out;
#pragma unroll
                        for (int j = 0; j < 16; ++j) amax = max(amax, abs(o[j]));
                        if (second) {
#pragma unroll
                            for (int j = 16; j < 32; ++j) amax = max(amax, abs(o[j]));
                        }
                        // (every thread writing its own row's 128 bytes straight from registers was measured: GEMM1 82.6 -> 104 us
                        // at 16K tokens, profiles/r3a_* — the staging transpose stays)
                        uint32_t blk[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) blk[j] = (uint32_t)o[j];
                        warp_store_block(st_addr, blk, obase + c * 4, pitch, rows_valid, 64, lane);
                        if (second) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) blk[j] = (uint32_t)o[16 + j];
                            warp_store_block(st_addr, blk, obase + (c + 16) * 4, pitch, rows_valid, 64, lane);
                        }
                    }
                }
                if (valid)
                    p.amax_part[w][(out_row0 + lane) * p.amax_ld[w] + tc.n_t * (kSplit ? 2 : 1) + (kSplit ? h : 0)] = amax;
            } else {
                const bool dequant = p.epi == I64B_EPI_DEQUANT;
                const float rs = (dequant && valid && p.row_scale[w] != nullptr) ? p.row_scale[w][a_row] : 1.0f;
                const bool f32out = p.out_dtype == I64B_F32;
                const int osz = f32out ? 4 : 2;
                uint8_t* obase = p.out[w] == nullptr ? nullptr
                                 : reinterpret_cast<uint8_t*>(p.out[w]) + (out_row0 * p.ld_out[w] + colw0) * osz;
                const int64_t pitch = p.ld_out[w] * osz;
#pragma unroll 1
                for (int c = 0; c < kColsW; c += 32) {
                    if (colw0 + c >= n_out) break;
                    const bool second = colw0 + c + 16 < n_out;
                    float4 yk0[4][kFuseMaxK], yk1[4][kFuseMaxK];       // routed rows of the chunk's two 16-column blocks
                    if (combine_tile && p.top_k > 0) {
                        const int rvl = (p.debug_flags & 4u) ? 0 : rows_valid;      // bring-up: bit2 = skip the routed-row loads
                        combine_issue_loads(cx, colw0 + c, rvl, lane, yk0);
                        combine_issue_loads(cx, colw0 + c + 16, second ? rvl : 0, lane, yk1);
                    }
                    uint32_t ar[32];
                    tmem_ld_32x32(taddr + c, ar);
                    tmem_ld_wait();
                    if constexpr (kInt8) {
                        if (p.raw_out != nullptr) {
                            uint8_t* rbase = reinterpret_cast<uint8_t*>(p.raw_out) + (out_row0 * p.ld_raw + colw0 + c) * 4;
                            uint32_t blk[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j) blk[j] = ar[j];
                            warp_store_block(st_addr, blk, rbase, p.ld_raw * 4, rows_valid, 64, lane);
                            if (second) {
#pragma unroll
                                for (int j = 0; j < 16; ++j) blk[j] = ar[16 + j];
                                warp_store_block(st_addr, blk, rbase + 64, p.ld_raw * 4, rows_valid, 64, lane);
                            }
                        }
                        if (dequant) {
#pragma unroll
                            for (int j4 = 0; j4 < 8; ++j4) {
                                const float4 s4 = ld_shared_f4(ws_addr + (c + j4 * 4) * 4);
                                // result_i32.float() * (x_scale * w_scale)              core/quantization.py:173
                                ar[4 * j4 + 0] = __float_as_uint(__fmul_rn((float)(int)ar[4 * j4 + 0], __fmul_rn(rs, s4.x)));
                                ar[4 * j4 + 1] = __float_as_uint(__fmul_rn((float)(int)ar[4 * j4 + 1], __fmul_rn(rs, s4.y)));
                                ar[4 * j4 + 2] = __float_as_uint(__fmul_rn((float)(int)ar[4 * j4 + 2], __fmul_rn(rs, s4.z)));
                                ar[4 * j4 + 3] = __float_as_uint(__fmul_rn((float)(int)ar[4 * j4 + 3], __fmul_rn(rs, s4.w)));
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) ar[j] = __float_as_uint((float)(int)ar[j]);
                        }
                    } else if (dequant) {   // bf16 GEMM with a per-column (and optional per-row) fp32 scale
#pragma unroll
                        for (int j4 = 0; j4 < 8; ++j4) {
                            const float4 s4 = ld_shared_f4(ws_addr + (c + j4 * 4) * 4);
                            ar[4 * j4 + 0] = __float_as_uint(__fmul_rn(__uint_as_float(ar[4 * j4 + 0]), __fmul_rn(rs, s4.x)));
                            ar[4 * j4 + 1] = __float_as_uint(__fmul_rn(__uint_as_float(ar[4 * j4 + 1]), __fmul_rn(rs, s4.y)));
                            ar[4 * j4 + 2] = __float_as_uint(__fmul_rn(__uint_as_float(ar[4 * j4 + 2]), __fmul_rn(rs, s4.z)));
                            ar[4 * j4 + 3] = __float_as_uint(__fmul_rn(__uint_as_float(ar[4 * j4 + 3]), __fmul_rn(rs, s4.w)));
                        }
                    }
                    if (combine_tile) {
                        const bool fo = p.final_dtype == I64B_F32;
                        const int fsz = fo ? 4 : 2;
                        const int64_t tok0 = (int64_t)tc.m_t * kBlockM + q * 32;
                        uint8_t* fbase = reinterpret_cast<uint8_t*>(p.final_out) + (tok0 * p.ld_final + colw0 + c) * fsz;
                        uint8_t* ykeep = (p.keep_shared_y && obase != nullptr) ? obase + c * 4 : nullptr;
                        uint32_t blk[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) blk[j] = ar[j];
                        warp_combine_block(st_addr, blk, cx, yk0, fbase, p.ld_final * fsz, fo, ykeep, pitch, rows_valid,
                                           lane);
                        if (second) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) blk[j] = ar[16 + j];
                            warp_combine_block(st_addr, blk, cx, yk1, fbase + 16 * fsz, p.ld_final * fsz, fo,
                                               ykeep ? ykeep + 64 : nullptr, pitch, rows_valid, lane);
                        }
                    } else if (obase != nullptr) {
                        uint32_t blk[16];
                        if (f32out) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) blk[j] = ar[j];
                            warp_store_block(st_addr, blk, obase + c * 4, pitch, rows_valid, 64, lane);
                            if (second) {
#pragma unroll
                                for (int j = 0; j < 16; ++j) blk[j] = ar[16 + j];
                                warp_store_block(st_addr, blk, obase + (c + 16) * 4, pitch, rows_valid, 64, lane);
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                blk[j] = pack_bf16x2(__uint_as_float(ar[2 * j]), __uint_as_float(ar[2 * j + 1]));
                            warp_store_block(st_addr, blk, obase + c * 2, pitch, rows_valid, second ? 64 : 32, lane);
                        }
                    }
                }
            }
            // accumulator stage drained: hand it back to the MMA thread
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (kCtas == 2) mbar_arrive_leader(&tempty_bar[as]); else mbar_arrive(&tempty_bar[as]);
            }
            if constexpr (!kDual) {
                if (p.fuse_combine && w == 0) {     // publish this warp's routed rows to the shared-expert tiles
                    if (!(p.debug_flags & 8u)) __threadfence();                     // bring-up: bit3 = no fence
                    __syncwarp();
                    if (lane == 0) red_release_gpu_add(p.tile_done, 1);
                }
            }
            if (tl && threadIdx.x == 64) tl[6] = clock64() - t0;
        }
    }
    if (tl && threadIdx.x == 0) tl[7] = clock64() - t0;

    tc_fence_before();
    if constexpr (kCtas == 2) cluster_sync_all(); else __syncthreads();    // nobody leaves while the peer still signals it
    if (warp == 1) {
        tc_fence_after();
        if constexpr (kCtas == 2) tmem_dealloc_cta2<kTmemCols>(tmem_base);
        else tmem_dealloc<kTmemCols>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Fused down projection + combine over the token-major layout (i64b_gemm.cuh, i64b_launch_gemm_tri).
// CTA pairs (M = 256 per unit), N = 256 output columns per tile (narrower tiles are bound by the SM's operand ingest: a
// 128-column version ran at 0.45 of the tensor peak), K in up to three segments — route 0, route 1, shared expert.
// TMEM holds two 256-column accumulators:
//     segment 0 -> acc0, segment 1 -> acc1
//     epilogue 1:  acc0 <- rs * (w0*acc0 + w1*acc1)      (tcgen05.ld / tcgen05.st; acc1 is released as soon as it is read)
//     segment 2 -> acc1  (the shared expert; runs while epilogue 1 still rewrites acc0)
//     epilogue 2:  out[token] = acc0 + ss*acc1
// i.e.  o = w0*y0;  o = o + w1*y1;  o = rs*o;  o = o + ss*y_shared  with the roundings of combine_kernel
// (layers/token_routed_mlp.py:156-185).  The fp32 per-assignment rows never leave the SM.
// ---------------------------------------------------------------------------------------------------------------
//
// kOneAcc: every route weight times the routed scale, and the shared scale, is a power of two (the named models: 0.5 / 0.5
// x 2.0, 1.0).  GEMM1's epilogue has then already multiplied the SwiGLU rows by those factors — exact in bf16 — and the
// block output is ONE sum over K = 2*I_e + S: the three segments accumulate into the same TMEM accumulator, the two
// 256-column accumulators alternate between tiles (tile i+1's MMAs run under tile i's epilogue, as in i64b_gemm_kernel),
// and the epilogue is a plain scattered store.  Same products, one fp32 accumulation instead of three partial sums.
//
// kInt8 (W8A8, three-accumulator form only): int8 operands, int32 accumulators; the epilogues dequantise every segment with
// its own per-row (re-quantised intermediate) and per-column (down-projection out-feature) scales,
// y = float(acc) * (s_row * s_col) as GEMM2's dequant epilogue does (core/quantization.py:173), then combine in the same
// order — bit-identical to the W8A8 chain (and through it to the reference arithmetic).  The folded routed sum goes back into acc0 as fp32 bits.
constexpr int kTriN = 256;
template <bool kOneAcc, int kType>
__global__ void __launch_bounds__(kThreads, 1)
i64b_gemm_tri_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB0,
                     const __grid_constant__ CUtensorMap tmB1, const __grid_constant__ I64bGemmParams p) {
    using Cfg = StageCfg<kTriN, 2>;
    constexpr int kStages = Cfg::kStages;
    constexpr int kStageB = Cfg::kStageB;              // this CTA's 128 weight rows
    constexpr bool kInt8 = kType == kTypeInt8;
    constexpr bool kFp8 = kType == kTypeFp8;           // E4M3 operands, fp32 accumulators, the same per-segment scales as W8A8
    constexpr bool kScaled = kInt8 || kFp8;
    constexpr int kBlockK = kScaled ? 128 : 64;        // elements per 128-byte row
    static_assert(!(kOneAcc && kScaled), "quantised segments carry their own scales: no shared accumulator");
    constexpr uint32_t kTmemCols = 512;                // acc0 | acc1
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + kStages * kStageA;
    uint8_t* misc = sB + kStages * kStageB;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(misc);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tfull_bar = empty_bar + kStages;         // [3]: segment s complete
    uint64_t* tempty_bar = tfull_bar + 4;              // [0]: acc1 read by epilogue 1;  [1]: both accumulators drained
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 4);
    int* prefix = reinterpret_cast<int*>(tmem_ptr + 2);
    I64bGroup* sg = reinterpret_cast<I64bGroup*>(misc + 512);
    uint8_t* stage_base = misc + kMiscBytes;
    static_assert((2 * 8 + 2 * 4) * 8 + 8 + 4 * (I64B_MAX_GROUPS + 1) <= 512, "barrier area");

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int unit0 = (int)blockIdx.x / 2;
    const int unit_step = (int)gridDim.x / 2;
    pdl_launch_dependents();

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB0);
        tma_prefetch_desc(&tmB1);
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 3; ++i) mbar_init(&tfull_bar[i], 1);
        for (int i = 0; i < 2; ++i) mbar_init(&tempty_bar[i], kEpiWarps * 2);
        fence_barrier_init();
    } else if (warp == 1) {
        tmem_alloc_cta2<kTmemCols>(tmem_ptr);
    }
    cluster_sync_all();
    pdl_wait_prior();

    const int ng = p.num_groups;
    for (int i = threadIdx.x; i < ng; i += kThreads) sg[i] = p.groups[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int g = 0; g < ng; ++g) {
            prefix[g] = acc;
            acc += ((sg[g].rows + 2 * kBlockM - 1) / (2 * kBlockM)) * ((sg[g].n_out + kTriN - 1) / kTriN);
        }
        prefix[ng] = acc;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const int total_tiles = prefix[ng];
    const int nkb_r = p.k[0] / kBlockK, nkb_s = p.k[1] / kBlockK;       // k-blocks per routed segment / of the shared one
    const int nseg = nkb_s > 0 ? 3 : 2;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0, g = 0;
            uint32_t phase = 0;
            for (int tile = unit0; tile < total_tiles; tile += unit_step) {
                TileCoord tc;
                decode_tile<kTriN, 2>(tile, g, prefix, sg, tc, rank);
                const I64bGroup grp = sg[tc.g];
                const int a_row = grp.row_start + tc.m_t * kBlockM;
                const int n0 = tc.n_t * kTriN + rank * (kTriN / 2);
                int a_col = 0;
                for (int seg = 0; seg < nseg; ++seg) {
                    const CUtensorMap* tb = seg == 2 ? &tmB1 : &tmB0;
                    const int b_row = (seg == 0 ? grp.w_row_base : (seg == 1 ? grp.pad0 : 0)) + n0;
                    const int nkb = seg == 2 ? nkb_s : nkb_r;
                    for (int kb = 0; kb < nkb; ++kb) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * (kStageA + kStageB));
                        tma_load_2d_cta2(sA + stage * kStageA, &tmA, &full_bar[stage], a_col + kb * kBlockK, a_row,
                                         I64B_L2_EVICT_NORMAL);
                        tma_load_2d_cta2(sB + stage * kStageB, tb, &full_bar[stage], kb * kBlockK, b_row,
                                         I64B_L2_EVICT_NORMAL);
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                    a_col += nkb * kBlockK;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = make_idesc<kType, kTriN, 2>();
            int stage = 0, iter = 0;
            uint32_t phase = 0;
            for (int tile = unit0; tile < total_tiles; tile += unit_step, ++iter) {
                const uint32_t par = (uint32_t)iter & 1u;
                if constexpr (kOneAcc) {
                    mbar_wait(&tempty_bar[par], (((uint32_t)iter >> 1) & 1u) ^ 1u);     // this accumulator's previous tile is drained
                    tc_fence_after();
                }
                for (int seg = 0; seg < nseg; ++seg) {
                    if constexpr (!kOneAcc) {
                        if (seg == 0) mbar_wait(&tempty_bar[1], par ^ 1);        // the previous tile's accumulators are drained
                        if (seg == 2) mbar_wait(&tempty_bar[0], par);            // epilogue 1 of THIS tile has read acc1
                        tc_fence_after();
                    }
                    const uint32_t d_tmem = tmem_base + (kOneAcc ? par * kTriN : (seg == 0 ? 0 : kTriN));
                    const int nkb = seg == 2 ? nkb_s : nkb_r;
                    for (int kb = 0; kb < nkb; ++kb) {
                        mbar_wait(&full_bar[stage], phase);
                        tc_fence_after();
                        const uint64_t a_desc = make_sw128_kmajor_desc(sA + stage * kStageA);
                        const uint64_t b_desc = make_sw128_kmajor_desc(sB + stage * kStageB);
                        const int fresh = kOneAcc ? (seg | kb) : kb;             // 0: the first MMA overwrites the accumulator
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            if constexpr (kInt8) umma_i8_cta2(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (fresh | k) ? 1u : 0u);
                            else if constexpr (kFp8) umma_f8_cta2(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (fresh | k) ? 1u : 0u);
                            else umma_bf16_cta2(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (fresh | k) ? 1u : 0u);
                        }
                        umma_commit_cta2(&empty_bar[stage]);
                        if (kb == nkb - 1) {
                            if constexpr (kOneAcc) {
                                if (seg == nseg - 1) umma_commit_cta2(&tfull_bar[par]);
                            } else {
                                umma_commit_cta2(&tfull_bar[seg]);
                            }
                        }
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else {
        const int q = warp & 3;
        const int h = (warp - 2) >> 2;
        const int r = q * 32 + lane;
        constexpr int kColsW = kTriN / 2;
        const int cbase = h * kColsW;
        const uint32_t st_addr = smem_u32(stage_base + (warp - 2) * kEpiWarpBytes);
        [[maybe_unused]] const uint32_t ws_addr = st_addr + kStageWarpBytes;
        const bool f32out = p.final_dtype == I64B_F32;
        const int fsz = f32out ? 4 : 2;
        const float w0 = p.w[0], w1 = p.w[1], rs = p.routed_scale, ss = p.shared_scale;
        uint8_t* obase = reinterpret_cast<uint8_t*>(p.final_out);
        const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + cbase;      // acc0, this warp's lanes and columns
        const uint32_t t1 = t0 + kTriN;                                           // acc1
        int g = 0, iter = 0;
        auto store_chunk = [&](const uint32_t (&v)[32], int c, int64_t my_off, int rows_valid) {
            uint32_t blk[16];
            if (f32out) {
#pragma unroll
                for (int j = 0; j < 16; ++j) blk[j] = v[j];
                warp_store_block_rows(st_addr, blk, obase + c * 4, my_off, rows_valid, 64, lane);
#pragma unroll
                for (int j = 0; j < 16; ++j) blk[j] = v[16 + j];
                warp_store_block_rows(st_addr, blk, obase + (c + 16) * 4, my_off, rows_valid, 64, lane);
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    blk[j] = pack_bf16x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
                warp_store_block_rows(st_addr, blk, obase + c * 2, my_off, rows_valid, 64, lane);
            }
        };
        for (int tile = unit0; tile < total_tiles; tile += unit_step, ++iter) {
            const uint32_t par = (uint32_t)iter & 1u;
            TileCoord tc;
            decode_tile<kTriN, 2>(tile, g, prefix, sg, tc, rank);
            const I64bGroup grp = sg[tc.g];
            const int row_in_group = tc.m_t * kBlockM + r;
            const bool valid = row_in_group < grp.rows;
            const int rows_valid = min(32, max(0, grp.rows - (tc.m_t * kBlockM + q * 32)));
            const int colw0 = tc.n_t * kTriN + cbase;
            int64_t my_off = 0;
            float wa = w0, wb = w1;         // weights of K segments 0 / 1 for this row (bit 30 of row_tok: route 1 sits in segment 0)
            [[maybe_unused]] float rsa = 0.0f, rsb = 0.0f, rsS = 0.0f;      // W8A8: this row's activation scales per segment
            if (valid) {
                const int rt = __ldg(p.row_tok + grp.row_start + row_in_group);
                my_off = ((int64_t)(rt & 0x3FFFFFFF) * p.ld_final + colw0) * fsz;
                if (rt & 0x40000000) { wa = w1; wb = w0; }
                if constexpr (kScaled) {
                    const float* sc = p.row_scale[0] + grp.row_start + row_in_group;     // [3, n_tokens] by pair-order row
                    rsa = __ldg(sc);
                    rsb = __ldg(sc + p.n_tokens);
                    rsS = __ldg(sc + 2 * (int64_t)p.n_tokens);
                }
            }
            if constexpr (kScaled) {
                // column scales of this warp's 128 columns: segments 0 / 1 into the (idle until epilogue 2) store staging
                // block, the shared expert's into the scale block
                __syncwarp();
#pragma unroll
                for (int i = 0; i < kColsW / 32; ++i) {
                    const int c = i * 32 + lane;
                    const bool cv = colw0 + c < grp.n_out;
                    st_f32(st_addr + c * 4, cv ? __ldg(p.col_scale[0] + grp.w_row_base + colw0 + c) : 0.0f);
                    st_f32(st_addr + (kColsW + c) * 4, cv ? __ldg(p.col_scale[0] + grp.pad0 + colw0 + c) : 0.0f);
                    st_f32(ws_addr + c * 4, cv ? __ldg(p.col_scale[1] + colw0 + c) : 0.0f);
                }
                __syncwarp();
            }
            if constexpr (kOneAcc) {
                mbar_wait(&tfull_bar[par], ((uint32_t)iter >> 1) & 1u);
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < kColsW; c += 32) {
                    uint32_t a0[32];
                    tmem_ld_32x32(t0 + par * kTriN + c, a0);
                    tmem_ld_wait();
                    store_chunk(a0, c, my_off, rows_valid);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(&tempty_bar[par]);
                continue;
            }
            mbar_wait(&tfull_bar[0], par);
            mbar_wait(&tfull_bar[1], par);
            tc_fence_after();
            // epilogue 1: the routed part; with a shared expert it goes back into acc0, else straight out
#pragma unroll 1
            for (int c = 0; c < kColsW; c += 32) {
                uint32_t a0[32], a1[32];
                tmem_ld_32x32(t0 + c, a0);
                tmem_ld_32x32(t1 + c, a1);
                tmem_ld_wait();
#pragma unroll
                if constexpr (kScaled) {
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 sa4 = ld_shared_f4(st_addr + (c + j4 * 4) * 4);
                        const float4 sb4 = ld_shared_f4(st_addr + (kColsW + c + j4 * 4) * 4);
                        const float sav[4] = {sa4.x, sa4.y, sa4.z, sa4.w};
                        const float sbv[4] = {sb4.x, sb4.y, sb4.z, sb4.w};
#pragma unroll
                        for (int jj = 0; jj < 4; ++jj) {
                            const int j = j4 * 4 + jj;
                            // result_i32.float() * (x_scale * w_scale)                      core/quantization.py:173
                            const float f0 = kInt8 ? (float)(int)a0[j] : __uint_as_float(a0[j]);
                            const float f1 = kInt8 ? (float)(int)a1[j] : __uint_as_float(a1[j]);
                            a0[j] = __float_as_uint(__fmul_rn(f0, __fmul_rn(rsa, sav[jj])));
                            a1[j] = __float_as_uint(__fmul_rn(f1, __fmul_rn(rsb, sbv[jj])));
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    float o = __fmul_rn(wa, __uint_as_float(a0[j]));          // a + b == b + a: same bits as w0*y0 + w1*y1
                    o = __fadd_rn(o, __fmul_rn(wb, __uint_as_float(a1[j])));
                    a0[j] = __float_as_uint(__fmul_rn(rs, o));
                }
                if (nseg == 3) tmem_st_32x32(t0 + c, a0);
                else store_chunk(a0, c, my_off, rows_valid);
            }
            if (nseg == 3) {
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(&tempty_bar[0]);       // acc1 may take the shared expert's segment
                mbar_wait(&tfull_bar[2], par);
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < kColsW; c += 32) {
                    uint32_t a0[32], a1[32];
                    tmem_ld_32x32(t0 + c, a0);
                    tmem_ld_32x32(t1 + c, a1);
                    tmem_ld_wait();
#pragma unroll
                    if constexpr (kScaled) {
#pragma unroll
                        for (int j4 = 0; j4 < 8; ++j4) {
                            const float4 s4 = ld_shared_f4(ws_addr + (c + j4 * 4) * 4);
                            const float sv[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                            for (int jj = 0; jj < 4; ++jj) {
                                const float f1 = kInt8 ? (float)(int)a1[j4 * 4 + jj] : __uint_as_float(a1[j4 * 4 + jj]);
                                a1[j4 * 4 + jj] = __float_as_uint(__fmul_rn(f1, __fmul_rn(rsS, sv[jj])));
                            }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        a0[j] = __float_as_uint(__fadd_rn(__uint_as_float(a0[j]), __fmul_rn(ss, __uint_as_float(a1[j]))));
                    store_chunk(a0, c, my_off, rows_valid);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&tempty_bar[1]);
        }
    }

    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_cta2<kTmemCols>(tmem_base);
    }
}

typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);

PFN_tmapEncodeTiled get_encode_fn() {
    static PFN_tmapEncodeTiled fn = nullptr;
    if (fn == nullptr) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_tmapEncodeTiled>(ptr);
    }
    return fn;
}

template <int kType, bool kDual, int kNMma, int kCtas, bool kGather = false>
int launch_inst(const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b0, const CUtensorMap& b1,
                const I64bGemmParams& p, int grid, cudaStream_t stream) {
    auto kern = i64b_gemm_kernel<kType, kDual, kNMma, kCtas, kGather>;
    constexpr int smem = StageCfg<kNMma, kCtas>::kSmem;
    static bool configured[I64B_MAX_DEVICES] = {};      // function attributes are per device (one table per instantiation)
    const int dev = i64b_current_device();
    if (!configured[dev]) {
        I64B_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured[dev] = true;
    }
    static const int dbg_flags = getenv("I64B_DEBUG_FLAGS") ? atoi(getenv("I64B_DEBUG_FLAGS")) : 0;
    static const bool dbg_tl = getenv("I64B_DEBUG_TIMELINE") != nullptr;
    if (kCtas == 2) {       // CTA pairs: cluster launch; the bring-up flags below are single-CTA only
        I64bGemmParams q = p;
        q.debug_flags = (uint32_t)dbg_flags & ~1u;
        I64B_CUDA(i64b_launch_cluster(kern, dim3(grid), dim3(kThreads), smem, stream, 2, a0, a1, b0, b1, q));
        i64b_count_launch();
        return I64B_OK;
    }
    if (dbg_flags == 0 && !dbg_tl) {
        I64B_CUDA(i64b_launch(kern, dim3(grid), dim3(kThreads), smem, stream, a0, a1, b0, b1, p));
        i64b_count_launch();
        return I64B_OK;
    }
    // ---- bring-up instrumentation only (allocates + synchronises; never active unless the env asks for it)
    I64bGemmParams q = p;
    q.debug_flags = (uint32_t)dbg_flags;
    static long long* d_tl = nullptr;
    if (dbg_tl && d_tl == nullptr) cudaMalloc(&d_tl, 256 * 8 * sizeof(long long));
    if (dbg_tl) { cudaMemsetAsync(d_tl, 0, 256 * 8 * sizeof(long long), stream); q.debug_timeline = d_tl; }
    kern<<<grid, kThreads, smem, stream>>>(a0, a1, b0, b1, q);
    I64B_LAUNCH_CHECK();
    i64b_count_launch();
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cs);
    if (dbg_tl && cs == cudaStreamCaptureStatusNone) {
        static int printed = 0;
        long long h[256 * 8];
        cudaStreamSynchronize(stream);
        cudaMemcpy(h, d_tl, sizeof(h), cudaMemcpyDeviceToHost);
        if (printed++ < 12) {
            long long g0 = h[0];
            for (int b = 0; b < grid; ++b) if (h[b * 8] && h[b * 8] < g0) g0 = h[b * 8];
            fprintf(stderr, "[i64b timeline] type=%d dual=%d nmma=%d grid=%d\n", kType, (int)kDual, kNMma, grid);
            for (int b = 0; b < grid; b += (grid > 16 ? grid / 8 : 1))
                fprintf(stderr, "  cta %3d start+%lldns prologue=%lld tma0=%lld full0=%lld lastmma=%lld epi0=%lld epiend=%lld end=%lld (cycles)\n",
                        b, h[b * 8] - g0, h[b * 8 + 1], h[b * 8 + 2], h[b * 8 + 3], h[b * 8 + 4], h[b * 8 + 5], h[b * 8 + 6], h[b * 8 + 7]);
        }
    }
    return I64B_OK;
}

}  // namespace

int i64b_current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= I64B_MAX_DEVICES) dev = 0;
    return dev;
}

int i64b_num_sms() {        // of the CURRENT device (a process may drive several)
    static int sms[I64B_MAX_DEVICES] = {};
    const int dev = i64b_current_device();
    if (sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sms[dev] = v;
    }
    return sms[dev];
}

int i64b_encode_tmap(CUtensorMap* out, const void* base, bool is_int8, int64_t rows, int64_t cols,
                     int64_t pitch_bytes, int box_rows) {
    PFN_tmapEncodeTiled fn = get_encode_fn();
    I64B_CHECK(fn != nullptr, I64B_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    I64B_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && pitch_bytes % 16 == 0, I64B_ERR_ARG,
               "TMA operand must be 16-byte aligned (base %p pitch %lld)", base, (long long)pitch_bytes);
    I64B_CHECK(rows > 0 && cols > 0 && box_rows > 0 && box_rows <= 256, I64B_ERR_ARG, "bad tensor-map shape");
    const int elt = is_int8 ? 1 : 2;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
    cuuint32_t box[2] = {(cuuint32_t)(kRowBytes / elt), (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, is_int8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                    const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    I64B_CHECK(r == CUDA_SUCCESS, I64B_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return I64B_OK;
}

// Upper bound on the CTAs of the GEMM launches made by the calling thread (0 = every SM).  The expert-parallel path leaves a
// few SMs to the NVLink exchange kernels that run next to the GEMMs on other streams (i64b_moe_set_cta_limit).
static thread_local int g_gemm_cta_limit = 0;
int i64b_set_gemm_cta_limit(int limit) {
    const int prev = g_gemm_cta_limit;
    g_gemm_cta_limit = limit > 0 ? limit : 0;
    return prev;
}

int i64b_launch_gemm(int type, bool dual, int n_mma, int ctas, const CUtensorMap& tmA0, const CUtensorMap& tmA1,
                     const CUtensorMap& tmB0, const CUtensorMap& tmB1, const I64bGemmParams& p,
                     int64_t max_tiles, cudaStream_t stream) {
    if (max_tiles <= 0) return I64B_OK;
    I64B_CHECK(type == kTypeBf16 || type == kTypeInt8 || type == kTypeFp8, I64B_ERR_ARG, "unknown operand type %d", type);
    I64B_CHECK(ctas == 1 || (ctas == 2 && n_mma == 256), I64B_ERR_ARG, "CTA pairs need the 256-wide tile");
    I64B_CHECK(p.num_groups >= 1 && p.num_groups <= I64B_MAX_GROUPS, I64B_ERR_ARG, "bad group count %d",
               p.num_groups);
    I64B_CHECK(p.k[0] > 0 && p.k[1] > 0, I64B_ERR_ARG, "K must be positive");
    I64B_CHECK(p.gather_idx == nullptr || ctas == 2, I64B_ERR_ARG, "row gather needs the pair kernel");
    int sms = i64b_num_sms();
    if (g_gemm_cta_limit > 0 && g_gemm_cta_limit < sms) sms = g_gemm_cta_limit >= 2 ? g_gemm_cta_limit : 2;
    if (ctas == 2) {        // max_tiles counts pair units here
        const int pairs = sms / 2;
        const int grid2 = 2 * (int)(max_tiles < pairs ? max_tiles : pairs);
        if (p.gather_idx != nullptr) {
            I64B_CHECK(type == kTypeBf16 && dual, I64B_ERR_ARG, "row gather is built for the bf16 GEMM1 pair kernel");
            return launch_inst<kTypeBf16, true, 256, 2, true>(tmA0, tmA1, tmB0, tmB1, p, grid2, stream);
        }
#define I64B_PAIR(T)                                                                                   \
        if (type == T && dual) return launch_inst<T, true, 256, 2>(tmA0, tmA1, tmB0, tmB1, p, grid2, stream);  \
        if (type == T) return launch_inst<T, false, 256, 2>(tmA0, tmA1, tmB0, tmB1, p, grid2, stream);
        I64B_PAIR(kTypeInt8) I64B_PAIR(kTypeFp8) I64B_PAIR(kTypeBf16)
#undef I64B_PAIR
    }
    const int grid = (int)(max_tiles < sms ? max_tiles : sms);
#define I64B_DISPATCH(T, DUAL, N)                                                               \
    if (type == T && dual == DUAL && n_mma == N)                                                \
        return launch_inst<T, DUAL, N, 1>(tmA0, tmA1, tmB0, tmB1, p, grid, stream);
#define I64B_DISPATCH_T(T)                                                                      \
    I64B_DISPATCH(T, false, 64) I64B_DISPATCH(T, false, 128) I64B_DISPATCH(T, false, 256)        \
    I64B_DISPATCH(T, true, 64) I64B_DISPATCH(T, true, 128) I64B_DISPATCH(T, true, 256)
    I64B_DISPATCH_T(kTypeBf16) I64B_DISPATCH_T(kTypeInt8) I64B_DISPATCH_T(kTypeFp8)
#undef I64B_DISPATCH_T
#undef I64B_DISPATCH
    i64b_set_error("unsupported GEMM tile configuration n_mma=%d", n_mma);
    return I64B_ERR_ARG;
}

int i64b_launch_gemm_tri(int type, bool one_acc, const CUtensorMap& tmA, const CUtensorMap& tmB0,
                         const CUtensorMap& tmB1, const I64bGemmParams& p, int64_t max_units, cudaStream_t stream) {
    if (max_units <= 0) return I64B_OK;
    const bool is_int8 = type != kTypeBf16;         // one-byte operands with per-segment scales (int8 or E4M3)
    I64B_CHECK(!(is_int8 && one_acc), I64B_ERR_ARG, "the single-accumulator form is bf16 only");
    I64B_CHECK(!is_int8 || (p.k[1] > 0 && p.k[0] % 128 == 0 && p.k[1] % 128 == 0 && p.row_scale[0] && p.col_scale[0] &&
                            p.col_scale[1]), I64B_ERR_ARG, "W8A8 fused down projection: shared expert, K %% 128 == 0 and scales");
    I64B_CHECK(p.groups != nullptr && p.num_groups >= 1 && p.num_groups <= I64B_MAX_GROUPS, I64B_ERR_ARG, "bad group count %d",
               p.num_groups);
    I64B_CHECK(p.k[0] > 0 && p.k[0] % 64 == 0 && p.k[1] >= 0 && p.k[1] % 64 == 0, I64B_ERR_ARG,
               "K segments must be multiples of 64 (got %d, %d)", p.k[0], p.k[1]);
    I64B_CHECK(p.row_tok && p.final_out, I64B_ERR_ARG, "null pointer");
    constexpr int smem = StageCfg<kTriN, 2>::kSmem;
    auto kern = type == kTypeInt8 ? i64b_gemm_tri_kernel<false, kTypeInt8>
                : (type == kTypeFp8 ? i64b_gemm_tri_kernel<false, kTypeFp8>
                   : (one_acc ? i64b_gemm_tri_kernel<true, kTypeBf16> : i64b_gemm_tri_kernel<false, kTypeBf16>));
    static bool configured[4][I64B_MAX_DEVICES] = {};
    const int dev = i64b_current_device();
    const int which = type == kTypeInt8 ? 2 : (type == kTypeFp8 ? 3 : (one_acc ? 1 : 0));
    if (!configured[which][dev]) {
        I64B_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured[which][dev] = true;
    }
    int sms = i64b_num_sms();
    if (g_gemm_cta_limit > 0 && g_gemm_cta_limit < sms) sms = g_gemm_cta_limit >= 2 ? g_gemm_cta_limit : 2;
    const int pairs = sms / 2;
    const int grid = 2 * (int)(max_units < pairs ? max_units : pairs);
    I64B_CUDA(i64b_launch_cluster(kern, dim3(grid), dim3(kThreads), smem, stream, 2, tmA, tmB0, tmB1, p));
    i64b_count_launch();
    return I64B_OK;
}
