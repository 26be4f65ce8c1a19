// NVLink probe (2 GPUs, one process): what do SM-issued peer accesses deliver, by access pattern?
// Rows of ROW bytes are moved between a local and a peer buffer; row order is sequential or a random permutation.
//   memcpy   : cudaMemcpyPeerAsync (copy engines) — the link's ceiling
//   st/ld    : one warp per row, 16-byte load/store-unit accesses (push / pull)
//   tma_st/ld: one thread per block streams rows through a shared-memory ring with cp.async.bulk (push / pull)
// "uni": GPU0 alone; "bi": both GPUs run the same thing toward each other at the same time.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/_build/nvlink_probe tools/nvlink_probe.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint64_t* b, uint32_t ph) {
    uint32_t ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(smem_u32(b)), "r"(ph) : "memory");
    return ok;
}
__device__ __forceinline__ void bulk_ld(void* s, const void* g, uint32_t bytes, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(smem_u32(s)), "l"(g), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void bulk_st(void* g, const void* s, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
}

// one warp per row: dst[perm_d[r]] = src[perm_s[r]]
template <int kUnroll>
__global__ void lsu_rows(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, const int* __restrict__ ps, const int* __restrict__ pd,
                         int rows, int row_bytes) {
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const int nvec = row_bytes / 16;
    for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows; r += warps) {
        const int4* s = reinterpret_cast<const int4*>(src + (size_t)ps[r] * row_bytes);
        int4* d = reinterpret_cast<int4*>(dst + (size_t)pd[r] * row_bytes);
        for (int i0 = 0; i0 < nvec; i0 += 32 * kUnroll) {
            int4 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) { const int i = i0 + u * 32 + lane; if (i < nvec) v[u] = s[i]; }
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) { const int i = i0 + u * 32 + lane; if (i < nvec) d[i] = v[u]; }
        }
    }
}

// one thread per block streams rows through a ring: bulk load src row -> smem, bulk store smem -> dst row
template <int kDepth>
__global__ void tma_rows(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, const int* __restrict__ ps, const int* __restrict__ pd,
                         int rows, int row_bytes) {
    extern __shared__ __align__(128) uint8_t ring[];
    __shared__ uint64_t full[kDepth];
    if (threadIdx.x != 0) return;
    for (int i = 0; i < kDepth; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const int per = (rows + gridDim.x - 1) / gridDim.x;
    const int r0 = blockIdx.x * per, cnt = max(0, min(rows, r0 + per) - r0);
    auto load = [&](int j) {
        const int slot = j % kDepth;
        mbar_expect(&full[slot], row_bytes);
        bulk_ld(ring + (size_t)slot * row_bytes, src + (size_t)ps[r0 + j] * row_bytes, row_bytes, &full[slot]);
    };
    for (int j = 0; j < min(cnt, kDepth - 1); ++j) load(j);
    for (int j = 0; j < cnt; ++j) {
        const int slot = j % kDepth;
        while (!mbar_try(&full[slot], (j / kDepth) & 1)) {}
        bulk_st(dst + (size_t)pd[r0 + j] * row_bytes, ring + (size_t)slot * row_bytes, row_bytes);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (j + kDepth - 1 < cnt) {
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            load(j + kDepth - 1);
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

struct Dev {
    int id;
    uint8_t *local, *recv;       // local source / local destination; recv is what the PEER writes into / reads from
    int *seq, *rnd;
    cudaStream_t s;
    cudaEvent_t a, b;
};

int main(int argc, char** argv) {
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t total = (size_t)64 << 20;                    // bytes moved per direction per test
    Dev d[2];
    for (int i = 0; i < 2; ++i) {
        d[i].id = i;
        CK(cudaSetDevice(i));
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, i, 1 - i));
        if (!can) { printf("no peer access\n"); return 0; }
        CK(cudaDeviceEnablePeerAccess(1 - i, 0));
        CK(cudaMalloc(&d[i].local, total));
        CK(cudaMalloc(&d[i].recv, total));
        CK(cudaMemset(d[i].local, i + 1, total));
        CK(cudaMemset(d[i].recv, 0, total));
        CK(cudaStreamCreate(&d[i].s));
        CK(cudaEventCreate(&d[i].a));
        CK(cudaEventCreate(&d[i].b));
        CK(cudaFuncSetAttribute(tma_rows<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CK(cudaFuncSetAttribute(tma_rows<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int row_sizes[] = {2048, 4096, 16384, 65536};
    printf("%-10s %-4s %-7s %-6s %-4s %9s %9s\n", "kind", "dir", "row", "order", "mode", "GB/s gpu0", "GB/s gpu1");
    for (int row_bytes : row_sizes) {
        const int rows = (int)(total / row_bytes);
        std::vector<int> seq(rows), rnd(rows);
        for (int i = 0; i < rows; ++i) seq[i] = rnd[i] = i;
        std::mt19937 g(7);
        std::shuffle(rnd.begin(), rnd.end(), g);
        for (int i = 0; i < 2; ++i) {
            CK(cudaSetDevice(i));
            CK(cudaMalloc(&d[i].seq, rows * 4));
            CK(cudaMalloc(&d[i].rnd, rows * 4));
            CK(cudaMemcpy(d[i].seq, seq.data(), rows * 4, cudaMemcpyHostToDevice));
            CK(cudaMemcpy(d[i].rnd, rnd.data(), rows * 4, cudaMemcpyHostToDevice));
        }
        // kind: 0 memcpy, 1 lsu, 2 tma(depth 8, 4 blocks/SM), 3 tma (depth 8, 1 block/SM), 4 lsu unroll 8
        for (int kind = 0; kind < 5; ++kind) {
            if (kind == 0 && row_bytes != 2048) continue;
            if ((kind == 2 || kind == 3) && row_bytes * 8 > 200 * 1024) continue;
            for (int pull = 0; pull < 2; ++pull) {
                for (int order = 0; order < 2; ++order) {         // 0: sequential rows on the remote side, 1: random
                    if (kind == 0 && (order || pull)) continue;
                    for (int bi = 0; bi < 2; ++bi) {
                        float ms[2] = {0, 0};
                        for (int rep = 0; rep < 3; ++rep) {
                            for (int i = 0; i < (bi ? 2 : 1); ++i) {
                                CK(cudaSetDevice(i));
                                const uint8_t* src = pull ? d[1 - i].recv : d[i].local;
                                uint8_t* dst = pull ? d[i].local : d[1 - i].recv;
                                const int* ps = pull ? (order ? d[i].rnd : d[i].seq) : d[i].seq;
                                const int* pd = pull ? d[i].seq : (order ? d[i].rnd : d[i].seq);
                                CK(cudaEventRecord(d[i].a, d[i].s));
                                if (kind == 0) CK(cudaMemcpyPeerAsync(dst, 1 - i, src, i, total, d[i].s));
                                else if (kind == 1) lsu_rows<4><<<sms * 8, 256, 0, d[i].s>>>(src, dst, ps, pd, rows, row_bytes);
                                else if (kind == 4) lsu_rows<8><<<sms * 8, 256, 0, d[i].s>>>(src, dst, ps, pd, rows, row_bytes);
                                else if (kind == 2) tma_rows<8><<<sms * (row_bytes <= 4096 ? 4 : 1), 32, row_bytes * 8, d[i].s>>>(src, dst, ps, pd, rows, row_bytes);
                                else tma_rows<4><<<sms, 32, row_bytes * 4, d[i].s>>>(src, dst, ps, pd, rows, row_bytes);
                                CK(cudaEventRecord(d[i].b, d[i].s));
                            }
                            for (int i = 0; i < (bi ? 2 : 1); ++i) {
                                CK(cudaSetDevice(i));
                                CK(cudaStreamSynchronize(d[i].s));
                                CK(cudaGetLastError());
                                CK(cudaEventElapsedTime(&ms[i], d[i].a, d[i].b));
                            }
                        }
                        const char* names[] = {"memcpy", "lsu_u4", "tma_d8x4", "tma_d4x1", "lsu_u8"};
                        printf("%-10s %-4s %-7d %-6s %-4s %9.1f %9.1f\n", names[kind], pull ? "pull" : "push", row_bytes,
                               order ? "random" : "seq", bi ? "bi" : "uni", total / ms[0] * 1e-6, bi ? total / ms[1] * 1e-6 : 0.0);
                    }
                }
            }
        }
        for (int i = 0; i < 2; ++i) { CK(cudaSetDevice(i)); CK(cudaFree(d[i].seq)); CK(cudaFree(d[i].rnd)); }
    }
    return 0;
}
