// Expert-parallel token exchange over NVLink / NVSwitch peer memory (BASELINE config 5).
//
// Reference: TokenRoutedMLP._forward_ep (layers/token_routed_i64/layer.py:209-298) — counts all-to-all + .tolist()
// host sync + stable sort by rank + two blocking all_to_all_single calls each way.  Here:
//   * routing is a pure function of (layer, token id), so a rank plans ONLY ITS OWN tokens with the single-GPU plan
//     kernels.  In that expert-major plan the experts of owner d form one contiguous slab; every receive buffer is cut
//     into one region per SOURCE rank, so the slab goes, row by row, into "my" region of d's buffer and nobody needs
//     anybody else's token ids: no id all-gather, no global plan, no split-size exchange, ranks may bring different
//     numbers of tokens;
//   * dispatch = one kernel that reads each local activation row once and writes it with 128-bit stores straight
//     into the OWNING rank's receive buffer (peer-mapped with CUDA IPC) — as bf16, or quantised once on the home rank
//     and sent as int8 + one fp32 scale per row (W8A8: half the bytes) — remembers, locally, (owner rank, row) of every
//     assignment, and publishes the per-expert row counts of each slab in the owner's flags page;
//   * the owner builds its GEMM group table (one group per (local expert, source rank)) from those counts;
//   * the return path is fused into the COMBINE: the home rank's combine kernel reads the K per-assignment rows of
//     each token straight from the owners' output buffers over NVLink (2-4 KB contiguous per row, 512 B per warp
//     instruction).  (Measured alternative, rejected: storing each output row to its home from the down-projection's
//     epilogue moves 64-byte segments over NVLink and ran at ~175 GB/s.);
//   * the only synchronisation is a flag barrier in peer memory (system-scope release / acquire, monotonically
//     increasing epochs kept on the device: graph-capturable, no host state).
#include <stdlib.h>

#include "i64b_gemm.cuh"
#include "i64b_plan.cuh"

namespace {

constexpr int kEpWarps = 4;
constexpr int kCountsBase = 64;       // flags page: ints [64 + src * 64 + local expert] = rows of that slab segment
constexpr int kCountsStride = 64;

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// One warp per LOCAL token: K peer-memory row copies + K (owner, row) tags.  Block 0 also publishes the slab counts.
template <bool kQuant>
__global__ void __launch_bounds__(kEpWarps * 32)
ep_dispatch_kernel(const i64b_ep_ctx c, const void* __restrict__ x, const I64bPlanWs ws, int64_t* __restrict__ loc) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int lane = threadIdx.x & 31;
    const int K = c.top_k, E = c.num_experts, bins = K * E, local_e = E / c.size;
    if (blockIdx.x == 0) {
        // rows this rank sends to expert e (all routes): a store into the owner's flags page; the barrier's system-scope
        // release publishes it together with the rows
        for (int e = threadIdx.x; e < E; e += blockDim.x) {
            const int rows = c.n_local > 0 ? ws.group_start[e + 1] - ws.group_start[e] : 0;
            c.peer_flags[e / local_e][kCountsBase + c.rank * kCountsStride + e % local_e] = rows;
        }
    }
    const int64_t t = (int64_t)blockIdx.x * kEpWarps + (threadIdx.x >> 5);
    if (t >= c.n_local) return;
    const int64_t n = c.n_local;
    int my_rank = 0, my_row = 0;
    if (lane < K) {
        const uint32_t pk = ws.packed[(int64_t)lane * n + t];
        const int e = pk & 0xFF;
        const int bin = lane * E + e;
        const int in_tile = (int)(pk >> 8) + ws.tile_counts[(t / I64B_PLAN_TILE) * bins + bin];
        const int dest = ws.group_start[e] + ws.route_base[bin] + in_tile;          // row in this rank's merged layout
        my_rank = e / local_e;
        my_row = (int)((int64_t)c.rank * c.cap_src) + dest - ws.group_start[my_rank * local_e];   // row in the owner's buffer
        loc[(int64_t)lane * n + t] = ((int64_t)my_rank << 32) | (int64_t)my_row;
    }
    const int H = c.hidden, nvec = H / 8;
    const int4* src = reinterpret_cast<const int4*>(reinterpret_cast<const __nv_bfloat16*>(x) + t * H);
    int dk[I64B_MAX_TOPK], rowk[I64B_MAX_TOPK];
#pragma unroll
    for (int k = 0; k < I64B_MAX_TOPK; ++k) {
        dk[k] = __shfl_sync(0xffffffffu, my_rank, k);
        rowk[k] = __shfl_sync(0xffffffffu, my_row, k);
    }
    if constexpr (!kQuant) {
        int4* dst[I64B_MAX_TOPK];
#pragma unroll
        for (int k = 0; k < I64B_MAX_TOPK; ++k)
            dst[k] = k < K ? reinterpret_cast<int4*>(reinterpret_cast<__nv_bfloat16*>(c.peer_xs[dk[k]]) + (int64_t)rowk[k] * H) : nullptr;
        // the whole row (H <= 2048: 8 vectors per lane) is requested before the first peer store leaves, so a warp pays one
        // local-memory latency per token, not one per 512 bytes
        constexpr int kV = 8;
        for (int i0 = 0; i0 < nvec; i0 += 32 * kV) {
            int4 v[kV];
#pragma unroll
            for (int j = 0; j < kV; ++j) {
                const int i = i0 + j * 32 + lane;
                if (i < nvec) v[j] = ld_nc_16(src + i);
            }
#pragma unroll
            for (int j = 0; j < kV; ++j) {
                const int i = i0 + j * 32 + lane;
                if (i < nvec) {
#pragma unroll
                    for (int k = 0; k < I64B_MAX_TOPK; ++k)
                        if (k < K) dst[k][i] = v[j];
                }
            }
        }
    } else {
        // quantize_activations_int8 (core/quantization.py:103-116) once per token; the K copies share row and scale
        float amax = 0.0f;
        for (int i = lane; i < nvec; i += 32) {
            const int4 v = __ldg(src + i);
            const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) amax = fmaxf(amax, fmaxf(fabsf(bf16lo(w[j])), fabsf(bf16hi(w[j]))));
        }
        amax = warp_max(amax);
        const float scale = quant_scale(amax);
        const float rcp = quant_rcp(scale);
        int8_t* dst[I64B_MAX_TOPK];
#pragma unroll
        for (int k = 0; k < I64B_MAX_TOPK; ++k)
            dst[k] = k < K ? reinterpret_cast<int8_t*>(c.peer_xs[dk[k]]) + (int64_t)rowk[k] * H : nullptr;
        for (int i = lane; i < nvec; i += 32) {
            const int4 v = __ldg(src + i);
            const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
            uint32_t pk[2] = {0, 0};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t q0 = (uint32_t)(quant_q_rcp(bf16lo(w[j]), scale, rcp) & 0xFF);
                const uint32_t q1 = (uint32_t)(quant_q_rcp(bf16hi(w[j]), scale, rcp) & 0xFF);
                pk[j >> 1] |= (q0 | (q1 << 8)) << ((j & 1) * 16);
            }
#pragma unroll
            for (int k = 0; k < I64B_MAX_TOPK; ++k)
                if (k < K) *reinterpret_cast<int2*>(dst[k] + i * 8) = make_int2((int)pk[0], (int)pk[1]);
        }
        if (lane < K) c.peer_xs_scale[my_rank][my_row] = scale;
    }
}

// GEMM group tables of THIS rank's experts over its receive buffer: one group per (local expert, source rank), built
// from the counts the sources published (one small block, after the barrier)
__global__ void ep_groups_kernel(const i64b_ep_ctx c, I64bGroup* g1, I64bGroup* g2, int expert_inter) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int local_e = c.num_experts / c.size;
    const int ng = local_e * c.size;
    const int* counts = c.peer_flags[c.rank] + kCountsBase;
    for (int i = threadIdx.x; i <= ng; i += blockDim.x) {
        if (i == ng) {      // no shared expert on this path (it runs on the home rank)
            g1[i] = I64bGroup{0, 0, 0, 16, 1, 0, 0, 0};
            g2[i] = I64bGroup{0, 0, 0, 16, 1, 0, 0, 0};
            continue;
        }
        const int le = i / c.size, r = i % c.size;
        int off = 0;
        for (int j = 0; j < le; ++j) off += counts[r * kCountsStride + j];
        const int rows = counts[r * kCountsStride + le];
        const int start = (int)((int64_t)r * c.cap_src) + off;
        g1[i] = I64bGroup{start, rows, le * 2 * expert_inter, expert_inter, 0, start, 0, 0};
        g2[i] = I64bGroup{start, rows, le * c.hidden, c.hidden, 0, start, 0, 0};
    }
}

// Combine on the home rank, pulling the per-assignment rows from their owners (same operation order as
// combine_kernel / layers/token_routed_mlp.py:156-185:  o = w0*y0; o = o + wk*yk; o = rs*o; o = o + ss*shared)
struct EpCombineArgs {
    const int64_t* loc;        // [K, n] (owner << 32) | row
    const float* y_shared;     // [n, H] fp32 or nullptr
    float w[I64B_MAX_TOPK];
    float routed_scale, shared_scale;
    void* out;
    int out_dtype;
};
struct Row8 { float v[8]; };
__device__ __forceinline__ void combine_accumulate(Row8& o, const Row8& y, float w, bool first) {
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = first ? __fmul_rn(w, y.v[j]) : __fadd_rn(o.v[j], __fmul_rn(w, y.v[j]));
}
// 8 consecutive columns of one peer row: one 16-byte load (bf16 transport) or two (fp32)
__device__ __forceinline__ void ep_issue_row_load(const i64b_ep_ctx& c, int64_t l, int col8, int4& a, int4& b) {
    const void* base = c.peer_y[(int)(l >> 32) & (I64B_EP_MAX_RANKS - 1)];
    const int64_t row = l & 0xffffffffll;
    if (c.y_dtype == I64B_F32) {
        const int4* p = reinterpret_cast<const int4*>(reinterpret_cast<const float*>(base) + row * c.hidden) + 2 * col8;
        a = ld_nc_16(p);
        b = ld_nc_16(p + 1);
    } else {
        a = ld_nc_16(reinterpret_cast<const int4*>(reinterpret_cast<const __nv_bfloat16*>(base) + row * c.hidden) + col8);
        b = make_int4(0, 0, 0, 0);
    }
}
__device__ __forceinline__ Row8 ep_unpack_row(int y_dtype, const int4& a, const int4& b) {
    Row8 r;
    if (y_dtype == I64B_F32) {
        r.v[0] = __int_as_float(a.x); r.v[1] = __int_as_float(a.y); r.v[2] = __int_as_float(a.z); r.v[3] = __int_as_float(a.w);
        r.v[4] = __int_as_float(b.x); r.v[5] = __int_as_float(b.y); r.v[6] = __int_as_float(b.z); r.v[7] = __int_as_float(b.w);
    } else {
        const uint32_t w[4] = {(uint32_t)a.x, (uint32_t)a.y, (uint32_t)a.z, (uint32_t)a.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) { r.v[2 * j] = bf16lo(w[j]); r.v[2 * j + 1] = bf16hi(w[j]); }
    }
    return r;
}
// One item = 8 output columns of one token.  A thread keeps kItems items in flight: all (owner, row) tags first, then all
// peer loads (K per item, 16 bytes each over NVLink), then the local shared-expert rows, then the arithmetic.
constexpr int kCombItems = 2;
__global__ void __launch_bounds__(256) ep_combine_kernel(const i64b_ep_ctx c, const EpCombineArgs a) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int h8 = c.hidden / 8;
    const int64_t total = (int64_t)c.n_local * h8;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += stride * kCombItems) {
        int64_t loc[kCombItems][I64B_MAX_TOPK];
        int4 ya[kCombItems][I64B_MAX_TOPK], yb[kCombItems][I64B_MAX_TOPK];
#pragma unroll
        for (int u = 0; u < kCombItems; ++u) {
            const int64_t i = i0 + u * stride;
            const int64_t t = i < total ? i / h8 : 0;
#pragma unroll
            for (int k = 0; k < I64B_MAX_TOPK; ++k)
                if (k < c.top_k) loc[u][k] = __ldg(a.loc + (int64_t)k * c.n_local + t);
        }
#pragma unroll
        for (int u = 0; u < kCombItems; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < total) {
                const int col8 = (int)(i % h8);
#pragma unroll
                for (int k = 0; k < I64B_MAX_TOPK; ++k)
                    if (k < c.top_k) ep_issue_row_load(c, loc[u][k], col8, ya[u][k], yb[u][k]);
            }
        }
#pragma unroll
        for (int u = 0; u < kCombItems; ++u) {
            const int64_t i = i0 + u * stride;
            if (i >= total) break;
            const int64_t t = i / h8;
            const int col8 = (int)(i % h8);
            Row8 o;
#pragma unroll
            for (int k = 0; k < I64B_MAX_TOPK; ++k)
                if (k < c.top_k) combine_accumulate(o, ep_unpack_row(c.y_dtype, ya[u][k], yb[u][k]), a.w[k], k == 0);
            const float rs = a.routed_scale;
#pragma unroll
            for (int j = 0; j < 8; ++j) o.v[j] = __fmul_rn(rs, o.v[j]);
            if (a.y_shared != nullptr) {
                const float4* sp = reinterpret_cast<const float4*>(a.y_shared + t * c.hidden) + 2 * col8;
                const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1);
                const float sv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
                const float ss = a.shared_scale;
#pragma unroll
                for (int j = 0; j < 8; ++j) o.v[j] = __fadd_rn(o.v[j], __fmul_rn(ss, sv[j]));
            }
            if (a.out_dtype == I64B_F32) {
                float4* op = reinterpret_cast<float4*>(a.out) + 2 * i;
                op[0] = make_float4(o.v[0], o.v[1], o.v[2], o.v[3]);
                op[1] = make_float4(o.v[4], o.v[5], o.v[6], o.v[7]);
            } else {
                uint32_t pk[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    __nv_bfloat162 v2 = __floats2bfloat162_rn(o.v[2 * j], o.v[2 * j + 1]);
                    pk[j] = *reinterpret_cast<uint32_t*>(&v2);
                }
                reinterpret_cast<int4*>(a.out)[i] = make_int4((int)pk[0], (int)pk[1], (int)pk[2], (int)pk[3]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Bulk-copy (TMA) versions of the two NVLink-facing kernels.  Measured with the LSU kernels above at 2 / 4 / 8 GPUs: the
// pull of the combine runs at 285-360 GB/s whatever the number of peers (16-byte loads from a peer take a full NVLink round
// trip, and an SM can only track a few dozen outstanding lines), a third to a half of the 770 GB/s the link delivers.
// cp.async.bulk moves a whole 2-4 KB row per instruction through the copy engine: no per-line bookkeeping in the SM, one
// thread keeps a ring of rows in flight, and the arithmetic reads the rows from shared memory.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int kPending> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(kPending) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait_trap(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity))
        if (clock64() - t0 > 20000000000ll) __trap();
}

constexpr int kTmaWarps = 8;            // independent streaming warps per block, each with its own ring and barriers
constexpr int kTmaMaxDepth = 8;
constexpr int kTmaSmemBudget = 200 * 1024;

// Dispatch, bf16 transport.  Every WARP owns a contiguous token range and a ring of `depth` rows in shared memory: lane 0
// streams the rows (bulk load row -> ring slot, then K bulk stores slot -> the owners' receive buffers over NVLink, the loads
// running depth-1 rows ahead); every 32 tokens all lanes decode the plan of the next 32 tokens into shared memory (and into
// `loc` for the combine).  No block-level synchronisation; a block is just kTmaWarps such warps sharing an SM.
__global__ void __launch_bounds__(kTmaWarps * 32)
ep_dispatch_tma_kernel(const i64b_ep_ctx c, const void* __restrict__ x, const I64bPlanWs ws, int64_t* __restrict__ loc,
                       int depth) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    __shared__ uint64_t full_bar[kTmaWarps][kTmaMaxDepth];
    __shared__ int64_t s_loc[kTmaWarps][32 * I64B_MAX_TOPK];
    pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int K = c.top_k, E = c.num_experts, bins = K * E, local_e = E / c.size;
    const uint32_t row_bytes = (uint32_t)c.hidden * 2;
    uint8_t* ring = smem_dyn + (size_t)warp * depth * row_bytes;
    uint64_t* full = full_bar[warp];
    if (lane == 0) {
        for (int i = 0; i < depth; ++i) mbar_init(&full[i], 1);
        fence_barrier_init();
    }
    __syncwarp();
    pdl_wait_prior();
    if (blockIdx.x == 0)
        for (int e = threadIdx.x; e < E; e += blockDim.x) {
            const int rows = c.n_local > 0 ? ws.group_start[e + 1] - ws.group_start[e] : 0;
            c.peer_flags[e / local_e][kCountsBase + c.rank * kCountsStride + e % local_e] = rows;
        }
    const int64_t n = c.n_local;
    const int64_t warps_total = (int64_t)gridDim.x * kTmaWarps;
    const int64_t per = (n + warps_total - 1) / warps_total;
    const int64_t t_begin = ((int64_t)blockIdx.x * kTmaWarps + warp) * per;
    const int total = (int)max((int64_t)0, min(n, t_begin + per) - t_begin);
    const uint8_t* xb = reinterpret_cast<const uint8_t*>(x) + t_begin * (int64_t)row_bytes;
    auto load_row = [&](int j) {                      // lane 0
        const int slot = j % depth;
        mbar_arrive_expect_tx(&full[slot], row_bytes);
        bulk_load(ring + (size_t)slot * row_bytes, xb + (int64_t)j * row_bytes, row_bytes, &full[slot]);
    };
    if (lane == 0)
        for (int j = 0; j < min(total, depth - 1); ++j) load_row(j);
    for (int j = 0; j < total; ++j) {
        if ((j & 31) == 0) {                          // the plan of the next 32 tokens -> (owner, row) tags
            __syncwarp();
            const int cnt = min(32, total - j);
            for (int i = lane; i < cnt * K; i += 32) {
                const int k = i / cnt, tt = i - k * cnt;
                const int64_t t = t_begin + j + tt;
                const uint32_t pk = ws.packed[(int64_t)k * n + t];
                const int e = pk & 0xFF;
                const int bin = k * E + e;
                const int in_tile = (int)(pk >> 8) + ws.tile_counts[(t / I64B_PLAN_TILE) * bins + bin];
                const int dest = ws.group_start[e] + ws.route_base[bin] + in_tile;
                const int owner = e / local_e;
                const int row = (int)((int64_t)c.rank * c.cap_src) + dest - ws.group_start[owner * local_e];
                const int64_t l = ((int64_t)owner << 32) | (int64_t)row;
                loc[(int64_t)k * n + t] = l;
                s_loc[warp][tt * K + k] = l;
            }
            __syncwarp();
        }
        if (lane == 0) {
            const int slot = j % depth;
            mbar_wait_trap(&full[slot], (uint32_t)(j / depth) & 1u);
            const uint8_t* src = ring + (size_t)slot * row_bytes;
            for (int k = 0; k < K; ++k) {
                const int64_t l = s_loc[warp][(j & 31) * K + k];
                bulk_store(reinterpret_cast<uint8_t*>(c.peer_xs[(int)(l >> 32) & (I64B_EP_MAX_RANKS - 1)]) +
                               (l & 0xffffffffll) * (int64_t)row_bytes, src, row_bytes);
            }
            bulk_commit();
            if (j + depth - 1 < total) {
                bulk_wait_read<1>();                  // the stores of row j-1 have read their slot: it takes row j + depth - 1
                load_row(j + depth - 1);
            }
        }
    }
    if (lane == 0) bulk_wait_all();
}

// Combine on the home rank.  Every WARP owns a contiguous token range and a ring of `depth` stages; lane 0 pulls, per token,
// the K per-assignment rows from their owners (one bulk copy per row over NVLink) and the local shared-expert row into a
// stage; the 32 lanes then apply the weights in the reference's operation order and write the output row.
__global__ void __launch_bounds__(kTmaWarps * 32)
ep_combine_tma_kernel(const i64b_ep_ctx c, const EpCombineArgs a, int depth, int stage_bytes) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    __shared__ uint64_t full_bar[kTmaWarps][kTmaMaxDepth];
    __shared__ int64_t s_loc[kTmaWarps][2][32 * I64B_MAX_TOPK];
    pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int K = c.top_k, H = c.hidden;
    const uint32_t ybytes = (uint32_t)H * (c.y_dtype == I64B_F32 ? 4u : 2u);
    const uint32_t sbytes = a.y_shared ? (uint32_t)H * 4u : 0u;
    uint8_t* ring = smem_dyn + (size_t)warp * depth * stage_bytes;
    uint64_t* full = full_bar[warp];
    if (lane == 0) {
        for (int i = 0; i < depth; ++i) mbar_init(&full[i], 1);
        fence_barrier_init();
    }
    __syncwarp();
    pdl_wait_prior();
    const int64_t n = c.n_local;
    const int64_t warps_total = (int64_t)gridDim.x * kTmaWarps;
    const int64_t per = (n + warps_total - 1) / warps_total;
    const int64_t t_begin = ((int64_t)blockIdx.x * kTmaWarps + warp) * per;
    const int total = (int)max((int64_t)0, min(n, t_begin + per) - t_begin);
    const int h8 = H / 8;
    auto load_tags = [&](int chunk) {                  // all lanes: (owner, row) tags of tokens [32*chunk, 32*chunk + 32)
        const int j0 = chunk * 32;
        const int cnt = min(32, total - j0);
        for (int i = lane; i < cnt * K; i += 32) {
            const int k = i / cnt, tt = i - k * cnt;
            s_loc[warp][chunk & 1][tt * K + k] = __ldg(a.loc + (int64_t)k * n + t_begin + j0 + tt);
        }
    };
    auto load_token = [&](int j) {                     // lane 0
        const int slot = j % depth;
        uint8_t* st = ring + (size_t)slot * stage_bytes;
        mbar_arrive_expect_tx(&full[slot], K * ybytes + sbytes);
        const int64_t* tags = s_loc[warp][(j >> 5) & 1] + (j & 31) * K;
        for (int k = 0; k < K; ++k) {
            const int64_t l = tags[k];
            bulk_load(st + (size_t)k * ybytes,
                      reinterpret_cast<const uint8_t*>(c.peer_y[(int)(l >> 32) & (I64B_EP_MAX_RANKS - 1)]) +
                          (l & 0xffffffffll) * (int64_t)ybytes, ybytes, &full[slot]);
        }
        if (sbytes) bulk_load(st + (size_t)K * ybytes, a.y_shared + (t_begin + j) * (int64_t)H, sbytes, &full[slot]);
    };
    if (total > 0) load_tags(0);
    if (total > 32) load_tags(1);
    __syncwarp();
    if (lane == 0)
        for (int j = 0; j < min(total, depth); ++j) load_token(j);
    for (int j = 0; j < total; ++j) {
        if ((j & 31) == 0 && j > 0 && j + 32 < total) {          // tokens ahead may belong to the next 32: fetch their tags now
            load_tags((j >> 5) + 1);
            __syncwarp();
        }
        const int slot = j % depth;
        mbar_wait_trap(&full[slot], (uint32_t)(j / depth) & 1u);
        const uint8_t* st = ring + (size_t)slot * stage_bytes;
        const int64_t t = t_begin + j;
        for (int c8 = lane; c8 < h8; c8 += 32) {
            Row8 o;
#pragma unroll
            for (int k = 0; k < I64B_MAX_TOPK; ++k) {
                if (k < K) {
                    const uint8_t* yr = st + (size_t)k * ybytes;
                    int4 ya, yb = make_int4(0, 0, 0, 0);
                    if (c.y_dtype == I64B_F32) {
                        ya = *reinterpret_cast<const int4*>(yr + (size_t)c8 * 32);
                        yb = *reinterpret_cast<const int4*>(yr + (size_t)c8 * 32 + 16);
                    } else {
                        ya = *reinterpret_cast<const int4*>(yr + (size_t)c8 * 16);
                    }
                    combine_accumulate(o, ep_unpack_row(c.y_dtype, ya, yb), a.w[k], k == 0);
                }
            }
            const float rs = a.routed_scale;
#pragma unroll
            for (int q = 0; q < 8; ++q) o.v[q] = __fmul_rn(rs, o.v[q]);
            if (sbytes) {
                const float4* sp = reinterpret_cast<const float4*>(st + (size_t)K * ybytes) + 2 * c8;
                const float4 s0 = sp[0], s1 = sp[1];
                const float sv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
                const float ss = a.shared_scale;
#pragma unroll
                for (int q = 0; q < 8; ++q) o.v[q] = __fadd_rn(o.v[q], __fmul_rn(ss, sv[q]));
            }
            const int64_t i = t * h8 + c8;
            if (a.out_dtype == I64B_F32) {
                float4* op = reinterpret_cast<float4*>(a.out) + 2 * i;
                op[0] = make_float4(o.v[0], o.v[1], o.v[2], o.v[3]);
                op[1] = make_float4(o.v[4], o.v[5], o.v[6], o.v[7]);
            } else {
                uint32_t pk[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    __nv_bfloat162 v2 = __floats2bfloat162_rn(o.v[2 * q], o.v[2 * q + 1]);
                    pk[q] = *reinterpret_cast<uint32_t*>(&v2);
                }
                reinterpret_cast<int4*>(a.out)[i] = make_int4((int)pk[0], (int)pk[1], (int)pk[2], (int)pk[3]);
            }
        }
        __syncwarp();                                  // every lane has read the stage: it takes token j + depth
        if (lane == 0 && j + depth < total) load_token(j + depth);
    }
}

// Flag barrier in peer memory.  signal: epoch = ++state[0]; flags[rank] of every peer := epoch (release.sys, after
// a system fence that orders this rank's earlier peer stores).  wait: until every flags[d] of this rank >= epoch.
// kMode: 1 signal, 2 wait, 3 both in one launch.
template <int kMode>
__global__ void ep_barrier_kernel(const i64b_ep_ctx c) {
    pdl_launch_dependents();
    pdl_wait_prior();
    __shared__ int epoch;
    if (threadIdx.x == 0) epoch = (kMode & 1) ? ++c.state[0] : c.state[0];
    __syncthreads();
    if ((int)threadIdx.x < c.size) {
        if (kMode & 1) {
            __threadfence_system();
            st_release_sys(c.peer_flags[threadIdx.x] + c.rank, epoch);
        }
        if (kMode & 2) {
            const long long t0 = clock64();
            while (ld_acquire_sys(c.peer_flags[c.rank] + threadIdx.x) < epoch) {
                __nanosleep(100);
                if (clock64() - t0 > 20000000000ll) __trap();       // ~10 s: a peer died; do not hang the GPU
            }
        }
    }
}

}  // namespace

static size_t ep_align(size_t v) { return (v + 255) & ~size_t(255); }
// I64B_EP_TMA=0 selects the load/store-unit versions of the dispatch and combine kernels (kept for comparison)
constexpr int kTmaSmemMax = kTmaSmemBudget;   // dynamic shared memory the bulk-copy kernels may ask for (rings of rows / stages)
static int ep_configure_tma() {           // function attributes are per device
    static bool configured[I64B_MAX_DEVICES] = {};
    const int dev = i64b_current_device();
    if (!configured[dev]) {
        I64B_CUDA(cudaFuncSetAttribute(ep_dispatch_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemMax));
        I64B_CUDA(cudaFuncSetAttribute(ep_combine_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemMax));
        configured[dev] = true;
    }
    return I64B_OK;
}
static bool ep_use_tma() {
    static const bool on = !(getenv("I64B_EP_TMA") && atoi(getenv("I64B_EP_TMA")) == 0);
    return on;
}

static int ep_check(const i64b_ep_ctx* c) {
    I64B_CHECK(c != nullptr, I64B_ERR_ARG, "null EP context");
    I64B_CHECK(c->size >= 1 && c->size <= I64B_EP_MAX_RANKS && c->rank >= 0 && c->rank < c->size, I64B_ERR_ARG,
               "EP rank %d / size %d out of range (max %d ranks)", c->rank, c->size, I64B_EP_MAX_RANKS);
    I64B_CHECK(c->num_experts >= 1 && c->num_experts <= I64B_MAX_EXPERTS && c->num_experts % c->size == 0, I64B_ERR_ARG,
               "num_experts (%d) must be divisible by ep_size (%d)", c->num_experts, c->size);
    I64B_CHECK(c->top_k >= 1 && c->top_k <= I64B_MAX_TOPK, I64B_ERR_ARG, "top_k %d out of range", c->top_k);
    I64B_CHECK(c->n_local >= 0 && c->hidden > 0 && c->hidden % 8 == 0, I64B_ERR_ARG, "bad EP geometry");
    I64B_CHECK(c->cap_src >= (int64_t)c->n_local * c->top_k && c->cap_src >= 1, I64B_ERR_ARG,
               "this rank's region holds %lld rows < %lld (n_local * top_k): reserve more tokens",
               (long long)c->cap_src, (long long)c->n_local * c->top_k);
    I64B_CHECK(c->cap_src * c->size < ((int64_t)1 << 31), I64B_ERR_ARG, "receive buffer too large for int32 rows");
    I64B_CHECK(c->x_mode == I64B_MODE_BF16 || c->x_mode == I64B_MODE_W8A8, I64B_ERR_ARG, "x_mode must be BF16 or W8A8");
    for (int d = 0; d < c->size; ++d)
        I64B_CHECK(c->peer_xs[d] && c->peer_y[d] && c->peer_flags[d] && (c->x_mode == I64B_MODE_BF16 || c->peer_xs_scale[d]),
                   I64B_ERR_ARG, "peer buffers of rank %d not mapped", d);
    I64B_CHECK(c->state != nullptr, I64B_ERR_ARG, "EP state missing");
    I64B_CHECK(c->y_dtype == I64B_F32 || c->y_dtype == I64B_BF16, I64B_ERR_ARG, "y_dtype must be F32 or BF16");
    return I64B_OK;
}

// EP workspace: local plan | (owner, row) of this rank's assignments | the owner-side GEMM group tables
struct EpWs {
    I64bPlanWs plan;
    int64_t* loc;
    I64bGroup *g1, *g2;
    size_t bytes;
};
static EpWs ep_carve(const i64b_ep_ctx* c, void* workspace) {
    EpWs w;
    uint8_t* base = workspace ? reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255))
                              : nullptr;
    const int64_t n_max = c->cap_src / c->top_k > 0 ? c->cap_src / c->top_k : 1;
    // the plan arrays are indexed with the call's n_local (packed is [K, n]); carving for n_max gives every array its
    // largest size, and the arrays after `packed` do not move with n
    w.plan = i64b_plan_ws_carve(base, n_max, c->top_k, c->num_experts);
    size_t off = ep_align(w.plan.bytes);
    w.loc = reinterpret_cast<int64_t*>(base ? base + off : nullptr);
    off += ep_align(sizeof(int64_t) * (size_t)c->top_k * (size_t)n_max);
    w.g1 = reinterpret_cast<I64bGroup*>(base ? base + off : nullptr);
    off += ep_align(sizeof(I64bGroup) * I64B_MAX_GROUPS);
    w.g2 = reinterpret_cast<I64bGroup*>(base ? base + off : nullptr);
    off += ep_align(sizeof(I64bGroup) * I64B_MAX_GROUPS);
    w.bytes = off;
    return w;
}

extern "C" {

int i64b_ep_alloc(size_t bytes, void** out) {
    I64B_CHECK(out != nullptr && bytes > 0, I64B_ERR_ARG, "bad allocation request");
    I64B_CUDA(cudaMalloc(out, bytes));
    I64B_CUDA(cudaMemset(*out, 0, bytes));
    return I64B_OK;
}
int i64b_ep_free(void* p) {
    if (p) I64B_CUDA(cudaFree(p));
    return I64B_OK;
}
int i64b_ipc_export(void* p, unsigned char* handle64) {
    I64B_CHECK(p && handle64, I64B_ERR_ARG, "null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    I64B_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle64, &h, 64);
    return I64B_OK;
}
int i64b_ipc_import(const unsigned char* handle64, void** out) {
    I64B_CHECK(handle64 && out, I64B_ERR_ARG, "null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    I64B_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return I64B_OK;
}
int i64b_ipc_close(void* p) {
    if (p) I64B_CUDA(cudaIpcCloseMemHandle(p));
    return I64B_OK;
}

size_t i64b_ep_workspace_bytes(const i64b_ep_ctx* c) {
    if (!c || c->top_k < 1 || c->cap_src < 1) return 0;
    return ep_carve(c, nullptr).bytes + 1024;
}

static int ep_run_plan(const i64b_ep_ctx* c, const int64_t* token_ids, const uint8_t* table_u8, const EpWs& w, cudaStream_t s) {
    I64bPlanArgs pa = {};
    pa.mode = I64B_ROUTE_TABLE;
    pa.token_ids = token_ids;
    pa.table = table_u8;
    pa.n = c->n_local;
    pa.vocab = c->vocab;
    pa.top_k = c->top_k;
    pa.num_experts = c->num_experts;
    pa.hidden = c->hidden;
    pa.ticket = c->state + 16;      // self-resetting: the last rank CTA finalises the plan in the same launch
    return i64b_run_plan(pa, w.plan, s);
}

int i64b_ep_plan(const i64b_ep_ctx* c, const int64_t* token_ids, const uint8_t* table_u8, void* workspace,
                 size_t workspace_bytes, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CHECK(table_u8 && workspace && (c->n_local == 0 || token_ids), I64B_ERR_ARG, "null pointer");
    I64B_CHECK(workspace_bytes >= i64b_ep_workspace_bytes(c), I64B_ERR_WORKSPACE, "EP plan workspace too small");
    if (c->n_local == 0) return I64B_OK;
    return ep_run_plan(c, token_ids, table_u8, ep_carve(c, workspace), (cudaStream_t)stream);
}

int i64b_ep_dispatch(const i64b_ep_ctx* c, const void* x, const int64_t* token_ids, const uint8_t* table_u8,
                     void* workspace, size_t workspace_bytes, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CHECK(table_u8 && workspace && (c->n_local == 0 || x), I64B_ERR_ARG, "null pointer");
    I64B_CHECK(workspace_bytes >= i64b_ep_workspace_bytes(c), I64B_ERR_WORKSPACE, "EP plan workspace too small");
    I64B_CHECK(c->num_experts / c->size <= kCountsStride, I64B_ERR_ARG, "too many local experts");
    const EpWs w = ep_carve(c, workspace);
    cudaStream_t s = (cudaStream_t)stream;
    if (c->n_local > 0 && token_ids != nullptr) {      // token_ids == NULL: the workspace holds i64b_ep_plan's plan
        rc = ep_run_plan(c, token_ids, table_u8, w, s);
        if (rc) return rc;
    }
    // a rank without tokens still publishes its (zero) counts: one block
    const int grid = (int)ceil_div64(c->n_local > 0 ? c->n_local : 1, kEpWarps);
    if (c->x_mode == I64B_MODE_W8A8) {
        I64B_CUDA(i64b_launch(ep_dispatch_kernel<true>, dim3(grid), dim3(kEpWarps * 32), 0, s, *c, x, w.plan, w.loc));
    } else if (ep_use_tma() && c->hidden % 8 == 0 && kTmaSmemBudget / (kTmaWarps * c->hidden * 2) >= 2) {
        if (int rc = ep_configure_tma()) return rc;
        // bulk-copy dispatch: one block of streaming warps per SM, or per SM of the exchange partition (ctx->comm_ctas)
        int depth = kTmaSmemBudget / (kTmaWarps * c->hidden * 2);
        if (depth > kTmaMaxDepth) depth = kTmaMaxDepth;
        const int blocks = c->comm_ctas > 0 ? c->comm_ctas : i64b_num_sms();
        I64B_CUDA(i64b_launch(ep_dispatch_tma_kernel, dim3(blocks), dim3(kTmaWarps * 32),
                              (size_t)kTmaWarps * depth * c->hidden * 2, s, *c, x, w.plan, w.loc, depth));
    } else {
        I64B_CUDA(i64b_launch(ep_dispatch_kernel<false>, dim3(grid), dim3(kEpWarps * 32), 0, s, *c, x, w.plan, w.loc));
    }
    i64b_count_launch(1);
    return I64B_OK;
}

int i64b_ep_combine(const i64b_ep_ctx* c, void* ep_workspace, const float* y_shared, const float* route_weight,
                    float routed_scale, float shared_scale, void* out, int out_dtype, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CHECK(ep_workspace && route_weight && (out || c->n_local == 0), I64B_ERR_ARG, "null pointer");
    I64B_CHECK(out_dtype == I64B_BF16 || out_dtype == I64B_F32, I64B_ERR_ARG, "unsupported output dtype");
    if (c->n_local == 0) return I64B_OK;
    EpCombineArgs a = {};
    a.loc = ep_carve(c, ep_workspace).loc;
    a.y_shared = y_shared;
    for (int k = 0; k < I64B_MAX_TOPK; ++k) a.w[k] = k < c->top_k ? route_weight[k] : 0.0f;
    a.routed_scale = routed_scale;
    a.shared_scale = shared_scale;
    a.out = out;
    a.out_dtype = out_dtype;
    {
        const int stage = c->top_k * c->hidden * (c->y_dtype == I64B_F32 ? 4 : 2) + (y_shared ? c->hidden * 4 : 0);
        int depth = kTmaSmemBudget / (kTmaWarps * stage);
        if (depth > 4) depth = 4;
        if (ep_use_tma() && depth >= 2) {
            if (int rc = ep_configure_tma()) return rc;
            const int blocks = c->comm_ctas > 0 ? c->comm_ctas : i64b_num_sms();
            I64B_CUDA(i64b_launch(ep_combine_tma_kernel, dim3(blocks), dim3(kTmaWarps * 32),
                                  (size_t)kTmaWarps * depth * stage, (cudaStream_t)stream, *c, a, depth, stage));
            i64b_count_launch(1);
            return I64B_OK;
        }
    }
    const int64_t total = (int64_t)c->n_local * (c->hidden / 8);
    const int64_t blocks = ceil_div64(total, 256 * kCombItems);
    const int64_t cap = (int64_t)i64b_num_sms() * 8;
    I64B_CUDA(i64b_launch(ep_combine_kernel, dim3((int)(blocks < cap ? blocks : cap)), dim3(256), 0, (cudaStream_t)stream,
                          *c, a));
    i64b_count_launch(1);
    return I64B_OK;
}

int i64b_ep_signal(const i64b_ep_ctx* c, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CUDA(i64b_launch(ep_barrier_kernel<1>, dim3(1), dim3(32), 0, (cudaStream_t)stream, *c));
    i64b_count_launch(1);
    return I64B_OK;
}
int i64b_ep_wait(const i64b_ep_ctx* c, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CUDA(i64b_launch(ep_barrier_kernel<2>, dim3(1), dim3(32), 0, (cudaStream_t)stream, *c));
    i64b_count_launch(1);
    return I64B_OK;
}
int i64b_ep_barrier(const i64b_ep_ctx* c, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CUDA(i64b_launch(ep_barrier_kernel<3>, dim3(1), dim3(32), 0, (cudaStream_t)stream, *c));
    i64b_count_launch(1);
    return I64B_OK;
}

}  // extern "C"

// owner side (called by i64b_moe_forward_dispatched in i64b_api.cu): group tables from the published counts
extern "C" int i64b_ep_build_groups(const i64b_ep_ctx* c, void* ep_workspace, int expert_inter, const I64bGroup** g1,
                                    const I64bGroup** g2, int* num_groups, cudaStream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    const EpWs w = ep_carve(c, ep_workspace);
    I64B_CUDA(i64b_launch(ep_groups_kernel, dim3(1), dim3(128), 0, stream, *c, w.g1, w.g2, expert_inter));
    i64b_count_launch(1);
    *g1 = w.g1;
    *g2 = w.g2;
    *num_groups = c->num_experts / c->size * c->size + 1;
    return I64B_OK;
}
