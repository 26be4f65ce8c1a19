// Bring-up probe (not part of the library): how fast can ONE launch per layer stream a decode layer's 33 MB of weights
// from HBM into shared memory with TMA bulk copies, as a function of ring depth, chunk size, CTAs/SM, PDL and an
// up-front L2 prefetch?  This is the ceiling for the fused decode kernel (csrc/i64b_fused.cu).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/stream_probe tools/stream_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.b32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(src), "r"(bytes) : "memory");
}

struct P {
    const uint8_t* w;      // this layer's weights
    size_t bytes;          // total bytes to stream
    int chunk;             // bytes per bulk copy
    int stages;
    int prefetch;          // 1: prefetch the CTA's whole share into L2 first
    int wait_frac;         // percent of the share that may be loaded before griddepcontrol.wait (0 = wait first)
    int* sink;
};

__global__ void __launch_bounds__(128, 2) stream_kernel(const __grid_constant__ P p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(base);
    uint8_t* ring = base + 1024;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const size_t nchunks = p.bytes / p.chunk;
    const size_t per = (nchunks + gridDim.x - 1) / gridDim.x;
    const size_t c0 = (size_t)blockIdx.x * per;
    const size_t c1 = c0 + per < nchunks ? c0 + per : nchunks;
    if (threadIdx.x == 0) {
        for (int i = 0; i < p.stages; ++i) mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (p.prefetch && threadIdx.x < 32) {
        for (size_t c = c0 + threadIdx.x; c < c1; c += 32) bulk_prefetch_l2(p.w + c * p.chunk, p.chunk);
    }
    if (threadIdx.x == 0) {
        const size_t n = c1 > c0 ? c1 - c0 : 0;
        const size_t early = n * p.wait_frac / 100;
        bool waited = false;
        size_t issued = 0, done = 0;
        int si = 0, sd = 0;
        uint32_t phase = 0;
        while (done < n) {
            while (issued < n && issued - done < (size_t)p.stages) {
                if (!waited && issued >= early) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }
                mbar_arrive_expect_tx(&full[si], p.chunk);
                bulk_load(ring + (size_t)si * p.chunk, p.w + (c0 + issued) * p.chunk, p.chunk, &full[si]);
                ++issued;
                if (++si == p.stages) si = 0;
            }
            while (!mbar_try_wait(&full[sd], phase)) { }
            ++done;
            if (++sd == p.stages) { sd = 0; phase ^= 1; }
        }
        if (!waited) asm volatile("griddepcontrol.wait;" ::: "memory");
        if (p.sink && ring[0] == 123 && ring[p.chunk] == 77) p.sink[0] = 1;
    }
}

__device__ __forceinline__ void tma_load_2d(void* dst, const void* desc, uint64_t* bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(smem_u32(dst)), "l"(desc), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// GEMM-like access: matrix [rows, 1024] bf16 (2 KB pitch); CTA b owns row blocks b, b+grid, ...; walks 16 k-blocks of
// (box_rows x 128 B) swizzled boxes per row block.
__global__ void __launch_bounds__(128, 2) tensor_kernel(const __grid_constant__ CUtensorMap tm, int rows, int box_rows,
                                                          int stages, int* sink) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(base);
    uint8_t* ring = base + 1024;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const int nblk = rows / box_rows;
        const int chunk = box_rows * 128;
        long n = 0;
        for (int b = blockIdx.x; b < nblk; b += gridDim.x) n += 16;
        long issued = 0, done = 0;
        int si = 0, sd = 0;
        uint32_t phase = 0;
        while (done < n) {
            while (issued < n && issued - done < stages) {
                const int blk = blockIdx.x + (int)(issued / 16) * gridDim.x, kb = (int)(issued % 16);
                mbar_arrive_expect_tx(&full[si], chunk);
                tma_load_2d(ring + (size_t)si * chunk, &tm, &full[si], kb * 64, blk * box_rows);
                ++issued;
                if (++si == stages) si = 0;
            }
            while (!mbar_try_wait(&full[sd], phase)) { }
            ++done;
            if (++sd == stages) { sd = 0; phase ^= 1; }
        }
        asm volatile("griddepcontrol.wait;" ::: "memory");
        if (sink && ring[0] == 123 && ring[chunk] == 77) sink[0] = 1;
    }
}

// GEMM-like access with a second, L2-resident operand: per k-block one (box_rows x 128 B) weight box from `tm` and
// a_rows x 128 B of the activation matrix `ta` (512 KB, shared by every CTA).
__global__ void __launch_bounds__(128, 1) ab_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap ta,
                                                      int rows, int a_rows, int stages, int* sink) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(base);
    uint8_t* ring = base + 1024;
    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const int nblk = rows / 128;
        const int chunk = 128 * 128 + a_rows * 128;
        long n = 0;
        for (int b = blockIdx.x; b < nblk; b += gridDim.x) n += 16;
        long issued = 0, done = 0;
        int si = 0, sd = 0;
        uint32_t phase = 0;
        while (done < n) {
            while (issued < n && issued - done < stages) {
                const int blk = blockIdx.x + (int)(issued / 16) * gridDim.x, kb = (int)(issued % 16);
                mbar_arrive_expect_tx(&full[si], chunk);
                tma_load_2d(ring + (size_t)si * 32768, &tm, &full[si], kb * 64, blk * 128);
                if (a_rows) tma_load_2d(ring + (size_t)si * 32768 + 16384, &ta, &full[si], kb * 64, (blk % 2) * 128);
                ++issued;
                if (++si == stages) si = 0;
            }
            while (!mbar_try_wait(&full[sd], phase)) { }
            ++done;
            if (++sd == stages) { sd = 0; phase ^= 1; }
        }
        if (sink && ring[0] == 123 && ring[chunk] == 77) sink[0] = 1;
    }
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                            const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

int main() {
    const int L = 24;
    const size_t bytes = 33030144;   // one tr-hash-moe-500m layer of bf16 expert + shared weights
    std::vector<uint8_t*> w(L);
    for (int i = 0; i < L; ++i) { CK(cudaMalloc(&w[i], bytes)); CK(cudaMemset(w[i], i + 1, bytes)); }
    int* sink; CK(cudaMalloc(&sink, 4));
    cudaStream_t s; CK(cudaStreamCreate(&s));
    CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    struct Cfg { int chunk, stages, occ, pdl, prefetch, wait_frac; };
    std::vector<Cfg> cfgs;
    for (int pdl = 0; pdl <= 1; ++pdl)
        for (int occ = 1; occ <= 2; ++occ)
            for (int pf = 0; pf <= 1; ++pf)
                for (int chunk : {8192, 16384, 32768}) {
                    const int budget = occ == 1 ? 196608 : 98304;
                    cfgs.push_back({chunk, budget / chunk, occ, pdl, pf, pdl ? 100 : 0});
                }
    cfgs.push_back({16384, 4, 1, 1, 1, 100});
    cfgs.push_back({16384, 4, 2, 1, 1, 100});
    cfgs.push_back({16384, 6, 2, 1, 1, 30});
    cfgs.push_back({16384, 6, 2, 1, 0, 30});
    for (const Cfg& c : cfgs) {
        const int smem = 2048 + c.chunk * c.stages;
        const int grid = 148 * c.occ;
        cudaGraph_t g; cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < L; ++i) {
            P p{w[i], bytes, c.chunk, c.stages, c.prefetch, c.wait_frac, sink};
            cudaLaunchConfig_t lc = {};
            lc.gridDim = dim3(grid); lc.blockDim = dim3(128); lc.dynamicSmemBytes = smem; lc.stream = s;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            lc.attrs = at; lc.numAttrs = c.pdl ? 1 : 0;
            CK(cudaLaunchKernelEx(&lc, stream_kernel, p));
        }
        CK(cudaStreamEndCapture(s, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        for (int i = 0; i < 5; ++i) CK(cudaGraphLaunch(ge, s));
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0, s));
        const int reps = 50;
        for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(ge, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double us = ms * 1e3 / (reps * L);
        printf("chunk=%5d stages=%2d occ=%d pdl=%d prefetch=%d early=%3d%% : %7.2f us/layer  %7.1f GB/s\n", c.chunk, c.stages,
               c.occ, c.pdl, c.prefetch, c.wait_frac, us, bytes / us / 1e3);
        CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
    }
    // ---- tensor-map (GEMM-like) streaming
    CK(cudaFuncSetAttribute(tensor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    void* fp = nullptr; cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr));
    PFN_enc enc = (PFN_enc)fp;
    const int rows = (int)(bytes / 2048);
    struct TC { int box_rows, stages, occ, grid_div, l2promo; };
    std::vector<TC> tcs = {{64, 12, 1, 1, 2}, {128, 6, 1, 1, 2}, {256, 3, 1, 1, 2}, {128, 10, 1, 1, 2}, {64, 6, 2, 1, 2}, {128, 3, 2, 1, 2},
                           {128, 6, 1, 1, 3}, {128, 6, 1, 1, 0}, {128, 6, 1, 2, 2}, {128, 6, 1, 4, 2}, {128, 6, 1, 16, 2}};
    for (const TC& c : tcs) {
        std::vector<CUtensorMap> tms(L);
        for (int i = 0; i < L; ++i) {
            cuuint64_t dims[2] = {1024, (cuuint64_t)rows};
            cuuint64_t strides[1] = {2048};
            cuuint32_t box[2] = {64, (cuuint32_t)c.box_rows};
            cuuint32_t es[2] = {1, 1};
            CUresult r = enc(&tms[i], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w[i], dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, (CUtensorMapL2promotion)c.l2promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
        }
        const int chunk = c.box_rows * 128;
        const int smem = 2048 + chunk * c.stages;
        const int grid = 148 * c.occ / c.grid_div;
        cudaGraph_t g; cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < L; ++i) {
            cudaLaunchConfig_t lc = {};
            lc.gridDim = dim3(grid); lc.blockDim = dim3(128); lc.dynamicSmemBytes = smem; lc.stream = s;
            CK(cudaLaunchKernelEx(&lc, tensor_kernel, tms[i], rows / c.grid_div, c.box_rows, c.stages, sink));
        }
        CK(cudaStreamEndCapture(s, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        for (int i = 0; i < 5; ++i) CK(cudaGraphLaunch(ge, s));
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0, s));
        const int reps = 50;
        for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(ge, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double us = ms * 1e3 / (reps * L);
        const double b = (double)bytes / c.grid_div;
        printf("tensor box_rows=%3d stages=%2d occ=%d grid=%3d l2promo=%d : %7.2f us/layer  %7.1f GB/s  (%.1f GB/s per CTA)\n", c.box_rows,
               c.stages, c.occ, grid, c.l2promo, us, b / us / 1e3, b / us / 1e3 / grid);
        CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
    }
    // ---- weight stream + L2-resident activation stream
    CK(cudaFuncSetAttribute(ab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    uint8_t* xa; CK(cudaMalloc(&xa, 256 * 2048)); CK(cudaMemset(xa, 3, 256 * 2048));
    struct AC { int a_rows, stages, grid, l2w; };
    std::vector<AC> acs = {{0, 6, 148, 0}, {64, 6, 148, 0}, {128, 6, 148, 0}, {0, 6, 104, 0}, {64, 6, 104, 0}, {128, 6, 104, 0},
                           {0, 6, 148, 1}, {64, 6, 148, 1}, {128, 6, 148, 1}, {0, 6, 104, 1}, {128, 6, 104, 1}, {128, 6, 32, 1}, {0, 6, 32, 1}};
    for (const AC& c : acs) {
        std::vector<CUtensorMap> tms(L);
        CUtensorMap ta;
        for (int i = 0; i <= L; ++i) {
            const bool isa = i == L;
            cuuint64_t dims[2] = {1024, (cuuint64_t)(isa ? 256 : rows)};
            cuuint64_t strides[1] = {2048};
            cuuint32_t box[2] = {64, (cuuint32_t)(isa ? (c.a_rows ? c.a_rows : 64) : 128)};
            cuuint32_t es[2] = {1, 1};
            CUresult r = enc(isa ? &ta : &tms[i], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, isa ? xa : w[c.l2w ? 0 : i], dims, strides, box, es,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
        }
        const int smem = 2048 + 32768 * c.stages;
        cudaGraph_t g; cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < L; ++i) {
            cudaLaunchConfig_t lc = {};
            lc.gridDim = dim3(c.grid); lc.blockDim = dim3(128); lc.dynamicSmemBytes = smem; lc.stream = s;
            CK(cudaLaunchKernelEx(&lc, ab_kernel, tms[i], ta, rows, c.a_rows, c.stages, sink));
        }
        CK(cudaStreamEndCapture(s, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        for (int i = 0; i < 5; ++i) CK(cudaGraphLaunch(ge, s));
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0, s));
        const int reps = 50;
        for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(ge, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double us = ms * 1e3 / (reps * L);
        const double tot = (double)bytes * (16384.0 + c.a_rows * 128.0) / 16384.0;
        printf("ab: a_rows=%3d grid=%3d weights_from=%s : %7.2f us/layer  W %7.1f GB/s  ingest %7.1f GB/s (%.1f GB/s per CTA)\n", c.a_rows, c.grid,
               c.l2w ? "L2  " : "DRAM", us, bytes / us / 1e3, tot / us / 1e3, tot / us / 1e3 / c.grid);
        CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
    }
    return 0;
}
