// Expert-parallel token exchange over NVLink / NVSwitch peer memory (BASELINE config 5).
//
// Reference: TokenRoutedMLP._forward_ep (layers/token_routed_i64/layer.py:209-298) — counts all-to-all + .tolist()
// host sync + stable sort by rank + two blocking all_to_all_single calls each way.  Here:
//   * routing is a pure function of (layer, token id), so after ONE small all-gather of the token ids every rank
//     computes the GLOBAL dispatch plan itself (the same plan kernels as on one GPU, over G*n tokens): no counts
//     exchange, no host round trip, no sort;
//   * dispatch = one kernel that reads each local activation row once and writes it with 128-bit stores straight
//     into the OWNING rank's expert-contiguous receive buffer (peer-mapped with CUDA IPC) and remembers, locally,
//     (owner rank, row) of every assignment;
//   * the return path is fused into the COMBINE: the home rank's combine kernel reads the K per-assignment fp32 rows
//     of each token straight from the owners' output buffers over NVLink (4 KB contiguous per row, 512 B per warp
//     instruction).  (Measured alternative, rejected: storing each output row to its home from the down-projection's
//     epilogue moves 64-byte segments over NVLink and ran at ~175 GB/s.);
//   * the only synchronisation is a flag barrier in peer memory (signal / wait kernels, system-scope release /
//     acquire, monotonically increasing epochs kept on the device: graph-capturable, no host state).
#include <stdlib.h>

#include "i64b_gemm.cuh"
#include "i64b_plan.cuh"

namespace {

constexpr int kEpWarps = 4;

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One warp per LOCAL token: K peer-memory row copies + K tags.  The plan was computed over all G*n tokens; this
// rank's token t is global token rank*n + t.
__global__ void __launch_bounds__(kEpWarps * 32)
ep_dispatch_kernel(const i64b_ep_ctx c, const void* __restrict__ x, const I64bPlanWs ws, int64_t* __restrict__ loc) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int lane = threadIdx.x & 31;
    const int64_t t = (int64_t)blockIdx.x * kEpWarps + (threadIdx.x >> 5);
    if (t >= c.n_local) return;
    const int K = c.top_k, E = c.num_experts, bins = K * E, local_e = E / c.size;
    const int64_t n_glob = (int64_t)c.n_local * c.size, gt = (int64_t)c.rank * c.n_local + t;
    int my_rank = 0, my_row = 0;
    if (lane < K) {
        const uint32_t pk = ws.packed[(int64_t)lane * n_glob + gt];
        const int e = pk & 0xFF;
        const int bin = lane * E + e;
        const int in_tile = (int)(pk >> 8) + ws.tile_counts[(gt / I64B_PLAN_TILE) * bins + bin];
        const int dest = ws.group_start[e] + ws.route_base[bin] + in_tile;      // row in the global merged layout
        my_rank = e / local_e;
        my_row = dest - ws.group_start[my_rank * local_e];                      // row in the owner's receive buffer
        loc[(int64_t)lane * c.n_local + t] = ((int64_t)my_rank << 32) | (int64_t)my_row;
    }
    const int H = c.hidden, nvec = H / 8;
    const int4* src = reinterpret_cast<const int4*>(reinterpret_cast<const __nv_bfloat16*>(x) + t * H);
    int4* dst[I64B_MAX_TOPK];
#pragma unroll
    for (int k = 0; k < I64B_MAX_TOPK; ++k) {
        const int d = __shfl_sync(0xffffffffu, my_rank, k), row = __shfl_sync(0xffffffffu, my_row, k);
        dst[k] = k < K ? reinterpret_cast<int4*>(reinterpret_cast<__nv_bfloat16*>(c.peer_xs[d]) + (int64_t)row * H) : nullptr;
    }
    for (int i = lane; i < nvec; i += 32) {
        const int4 v = ld_nc_16(src + i);
#pragma unroll
        for (int k = 0; k < I64B_MAX_TOPK; ++k)
            if (k < K) dst[k][i] = v;
    }
}

// GEMM group tables of THIS rank's experts in its receive buffer (one small block)
__global__ void ep_groups_kernel(const i64b_ep_ctx c, const I64bPlanWs ws, int expert_inter) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int local_e = c.num_experts / c.size, first = c.rank * local_e;
    const int le = threadIdx.x;
    if (le < local_e) {
        const int e = first + le;
        const int start = ws.group_start[e] - ws.group_start[first];
        const int rows = ws.group_start[e + 1] - ws.group_start[e];
        ws.groups1[le] = I64bGroup{start, rows, le * 2 * expert_inter, expert_inter, 0, start, 0, 0};
        ws.groups2[le] = I64bGroup{start, rows, le * c.hidden, c.hidden, 0, start, 0, 0};
    } else if (le == local_e) {      // no shared expert on this path (it runs replicated, on the home rank)
        ws.groups1[le] = I64bGroup{0, 0, 0, 16, 1, 0, 0, 0};
        ws.groups2[le] = I64bGroup{0, 0, 0, 16, 1, 0, 0, 0};
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
__global__ void __launch_bounds__(256) ep_combine_kernel(const i64b_ep_ctx c, const EpCombineArgs a) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int hv = c.hidden / 4;
    const int64_t total = (int64_t)c.n_local * hv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = i / hv;
        const int col = (int)(i % hv);
        float4 yk[I64B_MAX_TOPK];
#pragma unroll
        for (int k = 0; k < I64B_MAX_TOPK; ++k) {
            if (k < c.top_k) {
                const int64_t l = __ldg(a.loc + (int64_t)k * c.n_local + t);
                const void* base = c.peer_y[(int)(l >> 32) & (I64B_EP_MAX_RANKS - 1)];
                const int64_t row = l & 0xffffffffll;
                if (c.y_dtype == I64B_F32) {
                    const int4 v = ld_nc_16(reinterpret_cast<const int4*>(reinterpret_cast<const float*>(base) + row * c.hidden) + col);
                    yk[k] = make_float4(__int_as_float(v.x), __int_as_float(v.y), __int_as_float(v.z), __int_as_float(v.w));
                } else {
                    const uint2 v = *(reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(base) + row * c.hidden) + col);
                    yk[k] = make_float4(__uint_as_float(v.x << 16), __uint_as_float(v.x & 0xffff0000u),
                                        __uint_as_float(v.y << 16), __uint_as_float(v.y & 0xffff0000u));
                }
            }
        }
        float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < I64B_MAX_TOPK; ++k) {
            if (k < c.top_k) {
                const float w = a.w[k];
                const float4 y = yk[k];
                if (k == 0) {
                    o = make_float4(__fmul_rn(w, y.x), __fmul_rn(w, y.y), __fmul_rn(w, y.z), __fmul_rn(w, y.w));
                } else {
                    o.x = __fadd_rn(o.x, __fmul_rn(w, y.x));
                    o.y = __fadd_rn(o.y, __fmul_rn(w, y.y));
                    o.z = __fadd_rn(o.z, __fmul_rn(w, y.z));
                    o.w = __fadd_rn(o.w, __fmul_rn(w, y.w));
                }
            }
        }
        const float rs = a.routed_scale;
        o = make_float4(__fmul_rn(rs, o.x), __fmul_rn(rs, o.y), __fmul_rn(rs, o.z), __fmul_rn(rs, o.w));
        if (a.y_shared != nullptr) {
            const float4 y = __ldg(reinterpret_cast<const float4*>(a.y_shared + t * c.hidden) + col);
            const float ss = a.shared_scale;
            o.x = __fadd_rn(o.x, __fmul_rn(ss, y.x));
            o.y = __fadd_rn(o.y, __fmul_rn(ss, y.y));
            o.z = __fadd_rn(o.z, __fmul_rn(ss, y.z));
            o.w = __fadd_rn(o.w, __fmul_rn(ss, y.w));
        }
        if (a.out_dtype == I64B_F32) {
            reinterpret_cast<float4*>(a.out)[i] = o;
        } else {
            __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
            reinterpret_cast<uint2*>(a.out)[i] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
        }
    }
}

// Flag barrier in peer memory.  signal: epoch = ++state[0]; flags[rank] of every peer := epoch (release.sys, after
// a system fence that orders this rank's earlier peer stores).  wait: until every flags[d] of this rank >= epoch.
__global__ void ep_signal_kernel(const i64b_ep_ctx c) {
    pdl_wait_prior();
    __shared__ int epoch;
    if (threadIdx.x == 0) epoch = ++c.state[0];
    __syncthreads();
    if ((int)threadIdx.x < c.size) {
        __threadfence_system();
        st_release_sys(c.peer_flags[threadIdx.x] + c.rank, epoch);
    }
}
__global__ void ep_wait_kernel(const i64b_ep_ctx c) {
    pdl_wait_prior();
    if ((int)threadIdx.x < c.size) {
        const int epoch = c.state[0];
        const long long t0 = clock64();
        while (ld_acquire_sys(c.peer_flags[c.rank] + threadIdx.x) < epoch) {
            __nanosleep(100);
            if (clock64() - t0 > 20000000000ll) __trap();       // ~10 s: a peer died; do not hang the GPU
        }
    }
}

}  // namespace

static int ep_check(const i64b_ep_ctx* c) {
    I64B_CHECK(c != nullptr, I64B_ERR_ARG, "null EP context");
    I64B_CHECK(c->size >= 1 && c->size <= I64B_EP_MAX_RANKS && c->rank >= 0 && c->rank < c->size, I64B_ERR_ARG,
               "EP rank %d / size %d out of range (max %d ranks)", c->rank, c->size, I64B_EP_MAX_RANKS);
    I64B_CHECK(c->num_experts >= 1 && c->num_experts <= I64B_MAX_EXPERTS && c->num_experts % c->size == 0, I64B_ERR_ARG,
               "num_experts (%d) must be divisible by ep_size (%d)", c->num_experts, c->size);
    I64B_CHECK(c->top_k >= 1 && c->top_k <= I64B_MAX_TOPK, I64B_ERR_ARG, "top_k %d out of range", c->top_k);
    I64B_CHECK(c->n_local >= 0 && c->hidden > 0 && c->hidden % 8 == 0, I64B_ERR_ARG, "bad EP geometry");
    I64B_CHECK(c->cap_rows >= (int64_t)c->n_local * c->size * c->top_k, I64B_ERR_ARG,
               "receive capacity %lld rows < worst case %lld", (long long)c->cap_rows,
               (long long)c->n_local * c->size * c->top_k);
    for (int d = 0; d < c->size; ++d)
        I64B_CHECK(c->peer_xs[d] && c->peer_y[d] && c->peer_flags[d], I64B_ERR_ARG,
                   "peer buffers of rank %d not mapped", d);
    I64B_CHECK(c->state != nullptr, I64B_ERR_ARG, "EP state missing");
    I64B_CHECK(c->y_dtype == I64B_F32 || c->y_dtype == I64B_BF16, I64B_ERR_ARG, "y_dtype must be F32 or BF16");
    return I64B_OK;
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
    if (!c) return 0;
    return i64b_plan_ws_carve(nullptr, (int64_t)c->n_local * c->size, c->top_k, c->num_experts).bytes + 1024 +
           sizeof(int64_t) * (size_t)c->top_k * (size_t)(c->n_local > 0 ? c->n_local : 1);
}
// (owner, row) table of this rank's assignments, behind the plan in the EP workspace
static int64_t* ep_loc(const i64b_ep_ctx* c, void* workspace) {
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
    const size_t plan = i64b_plan_ws_carve(nullptr, (int64_t)c->n_local * c->size, c->top_k, c->num_experts).bytes;
    return reinterpret_cast<int64_t*>(base + ((plan + 255) & ~size_t(255)));
}

int i64b_ep_dispatch(const i64b_ep_ctx* c, const void* x, const int64_t* ids_all, const uint8_t* table_u8,
                     int expert_inter, void* workspace, size_t workspace_bytes, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CHECK(x && ids_all && table_u8 && workspace, I64B_ERR_ARG, "null pointer");
    I64B_CHECK(workspace_bytes >= i64b_ep_workspace_bytes(c), I64B_ERR_WORKSPACE, "EP plan workspace too small");
    if (c->n_local == 0) return I64B_OK;
    void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
    const int64_t n_glob = (int64_t)c->n_local * c->size;
    const I64bPlanWs ws = i64b_plan_ws_carve(base, n_glob, c->top_k, c->num_experts);
    I64bPlanArgs pa = {};
    pa.mode = I64B_ROUTE_TABLE;
    pa.token_ids = ids_all;
    pa.table = table_u8;
    pa.n = n_glob;
    pa.vocab = c->vocab;
    pa.top_k = c->top_k;
    pa.num_experts = c->num_experts;
    pa.expert_inter = expert_inter;
    pa.hidden = c->hidden;
    rc = i64b_run_plan(pa, ws, (cudaStream_t)stream);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    I64B_CUDA(i64b_launch(ep_dispatch_kernel, dim3((int)ceil_div64(c->n_local, kEpWarps)), dim3(kEpWarps * 32), 0, s, *c,
                          x, ws, ep_loc(c, workspace)));
    I64B_CUDA(i64b_launch(ep_groups_kernel, dim3(1), dim3(I64B_MAX_EXPERTS + 32), 0, s, *c, ws, expert_inter));
    i64b_count_launch(2);
    return I64B_OK;
}

int i64b_ep_combine(const i64b_ep_ctx* c, void* ep_workspace, const float* y_shared, const float* route_weight,
                    float routed_scale, float shared_scale, void* out, int out_dtype, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CHECK(ep_workspace && route_weight && out, I64B_ERR_ARG, "null pointer");
    I64B_CHECK(out_dtype == I64B_BF16 || out_dtype == I64B_F32, I64B_ERR_ARG, "unsupported output dtype");
    I64B_CHECK(c->hidden % 4 == 0, I64B_ERR_ARG, "hidden size must be a multiple of 4");
    if (c->n_local == 0) return I64B_OK;
    EpCombineArgs a = {};
    a.loc = ep_loc(c, ep_workspace);
    a.y_shared = y_shared;
    for (int k = 0; k < I64B_MAX_TOPK; ++k) a.w[k] = k < c->top_k ? route_weight[k] : 0.0f;
    a.routed_scale = routed_scale;
    a.shared_scale = shared_scale;
    a.out = out;
    a.out_dtype = out_dtype;
    const int64_t total = (int64_t)c->n_local * (c->hidden / 4);
    const int64_t blocks = ceil_div64(total, 256);
    const int64_t cap = (int64_t)i64b_num_sms() * 8;
    I64B_CUDA(i64b_launch(ep_combine_kernel, dim3((int)(blocks < cap ? blocks : cap)), dim3(256), 0, (cudaStream_t)stream,
                          *c, a));
    i64b_count_launch(1);
    return I64B_OK;
}

int i64b_ep_signal(const i64b_ep_ctx* c, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CUDA(i64b_launch(ep_signal_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, *c));
    i64b_count_launch(1);
    return I64B_OK;
}
int i64b_ep_wait(const i64b_ep_ctx* c, i64b_stream_t stream) {
    int rc = ep_check(c);
    if (rc) return rc;
    I64B_CUDA(i64b_launch(ep_wait_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, *c));
    i64b_count_launch(1);
    return I64B_OK;
}

}  // extern "C"

// group tables written by i64b_ep_dispatch (for i64b_moe_forward_dispatched in i64b_api.cu)
extern "C" void i64b_ep_group_tables(const i64b_ep_ctx* c, void* workspace, const I64bGroup** g1, const I64bGroup** g2) {
    void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
    const I64bPlanWs ws = i64b_plan_ws_carve(base, (int64_t)c->n_local * c->size, c->top_k, c->num_experts);
    *g1 = ws.groups1;
    *g2 = ws.groups2;
}
