// The step either side of the block: fused residual-add + RMSNorm (+ per-token INT8 quantisation emitting
// x_preq) in ONE pass over the hidden state.
//
// Reference call site  models/complexity_deep/model.py:352-363:
//     hidden = residual + hidden            (attention output added to the residual stream, model dtype)
//     residual = hidden
//     hidden = self.post_attention_layernorm(hidden)          layers/rmsnorm.py:83-90
//     hidden = self.mlp(hidden, token_ids=token_ids)          (int8 MLPs take x_preq = (int8, scale): layers/dense_mlp.py:46-58)
// and the reference's own fused kernel  csrc/I64_rmsnorm.cu:98-190 (I64_rmsnorm_quant_kernel: fp32 in, three passes over
// the row, one block per row).  Here one warp owns a row, reads it ONCE with 128-bit loads (the row stays in registers
// for H <= 2048), and writes the new residual, the normalised row and/or its int8 form + scale: per row 2·2H bytes in,
// 2H (+2H) (+H+4) out instead of the five separate elementwise kernels and four [N,H] round trips of the eager path.
//
// Arithmetic (bf16 model dtype; the eager PyTorch path of layers/rmsnorm.py:86-90 defines it):
//     h   = bf16(residual + x)                                   bf16 + bf16 in fp32, one rounding
//     r   = rsqrt(mean(h_f32^2) + eps)                           fp32
//     n1  = bf16(h_f32 * r)                                      "(x.float() * norm).type_as(x)"
//     y   = bf16(n1 * w)                                         bf16 * bf16 weight, one rounding
//     s   = max(absmax(y), 1e-8) / 127 ; q = clamp(rint(y / s))  quantize_activations_int8, core/quantization.py:103-116
// fp32 inputs (the csrc kernel's contract): y = (x * r) * w in fp32, no intermediate rounding (I64_rmsnorm.cu:147).
#include "i64b_common.cuh"

void i64b_count_launch(int n);

namespace {

constexpr int kWarps = 4;
constexpr int kMaxVec = 8;          // 16-byte vectors per lane kept in registers: H <= 32 * 8 * 8 = 2048 (bf16)

__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float rbf(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

struct NormArgs {
    const void* x;            // [N, H] bf16 | f32
    const void* residual;     // [N, H] same dtype, or nullptr
    const void* weight;       // [H] same dtype as x
    void* h_out;              // [N, H] new residual (residual + x), or nullptr
    void* y_out;              // [N, H] normalised rows, or nullptr
    int8_t* q_out;            // [N, H] int8, or nullptr
    float* scale_out;         // [N]
    int64_t n;
    int hidden;
    float eps;
};

// bf16: one warp per row, the row lives in registers as packed bf16 pairs
__global__ void __launch_bounds__(kWarps * 32) add_rmsnorm_bf16_kernel(const NormArgs a) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (row >= a.n) return;
    const int H = a.hidden, nvec = H / 8;
    const int4* xs = reinterpret_cast<const int4*>(reinterpret_cast<const __nv_bfloat16*>(a.x) + row * H);
    const int4* rs = a.residual ? reinterpret_cast<const int4*>(reinterpret_cast<const __nv_bfloat16*>(a.residual) + row * H)
                                : nullptr;
    const int4* ws = reinterpret_cast<const int4*>(a.weight);
    uint32_t hreg[kMaxVec][4];
    float ss = 0.0f;
#pragma unroll
    for (int j = 0; j < kMaxVec; ++j) {
        const int i = j * 32 + lane;
        if (i < nvec) {
            const int4 xv = ld_nc_16(xs + i);
            uint32_t w[4] = {(uint32_t)xv.x, (uint32_t)xv.y, (uint32_t)xv.z, (uint32_t)xv.w};
            if (rs != nullptr) {
                const int4 rv = ld_nc_16(rs + i);
                const uint32_t r[4] = {(uint32_t)rv.x, (uint32_t)rv.y, (uint32_t)rv.z, (uint32_t)rv.w};
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    w[c] = pack2(__fadd_rn(bf16lo(r[c]), bf16lo(w[c])), __fadd_rn(bf16hi(r[c]), bf16hi(w[c])));
                if (a.h_out != nullptr)
                    st_na_16(reinterpret_cast<int4*>(reinterpret_cast<__nv_bfloat16*>(a.h_out) + row * H) + i,
                             make_int4((int)w[0], (int)w[1], (int)w[2], (int)w[3]));
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                hreg[j][c] = w[c];
                const float lo = bf16lo(w[c]), hi = bf16hi(w[c]);
                ss = __fmaf_rn(lo, lo, ss);
                ss = __fmaf_rn(hi, hi, ss);
            }
        }
    }
    ss = warp_sum_f(ss);
    const float r = rsqrtf(__fadd_rn(__fdiv_rn(ss, (float)H), a.eps));
    float amax = 0.0f;
#pragma unroll
    for (int j = 0; j < kMaxVec; ++j) {
        const int i = j * 32 + lane;
        if (i < nvec) {
            const int4 wv = __ldg(ws + i);
            const uint32_t w[4] = {(uint32_t)wv.x, (uint32_t)wv.y, (uint32_t)wv.z, (uint32_t)wv.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float lo = rbf(__fmul_rn(rbf(__fmul_rn(bf16lo(hreg[j][c]), r)), bf16lo(w[c])));
                const float hi = rbf(__fmul_rn(rbf(__fmul_rn(bf16hi(hreg[j][c]), r)), bf16hi(w[c])));
                hreg[j][c] = pack2(lo, hi);
                amax = fmaxf(amax, fmaxf(fabsf(lo), fabsf(hi)));
            }
            if (a.y_out != nullptr)
                st_na_16(reinterpret_cast<int4*>(reinterpret_cast<__nv_bfloat16*>(a.y_out) + row * H) + i,
                         make_int4((int)hreg[j][0], (int)hreg[j][1], (int)hreg[j][2], (int)hreg[j][3]));
        }
    }
    if (a.q_out == nullptr) return;
    amax = warp_max(amax);
    const float scale = quant_scale(amax);
    const float rcp = quant_rcp(scale);
#pragma unroll
    for (int j = 0; j < kMaxVec; ++j) {
        const int i = j * 32 + lane;
        if (i < nvec) {
            uint32_t pk[2] = {0, 0};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint32_t q0 = (uint32_t)(quant_q_rcp(bf16lo(hreg[j][c]), scale, rcp) & 0xFF);
                const uint32_t q1 = (uint32_t)(quant_q_rcp(bf16hi(hreg[j][c]), scale, rcp) & 0xFF);
                pk[c >> 1] |= (q0 | (q1 << 8)) << ((c & 1) * 16);
            }
            st_na_8(a.q_out + row * H + (int64_t)i * 8, make_int2((int)pk[0], (int)pk[1]));
        }
    }
    if (lane == 0) a.scale_out[row] = scale;
}

// fp32 (the contract of csrc/I64_rmsnorm.cu): one warp per row, the row is re-read from L1/L2 for the later passes
__global__ void __launch_bounds__(kWarps * 32) add_rmsnorm_f32_kernel(const NormArgs a) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (row >= a.n) return;
    const int H = a.hidden;
    const float* x = reinterpret_cast<const float*>(a.x) + row * H;
    const float* res = a.residual ? reinterpret_cast<const float*>(a.residual) + row * H : nullptr;
    const float* w = reinterpret_cast<const float*>(a.weight);
    float* hout = a.h_out ? reinterpret_cast<float*>(a.h_out) + row * H : nullptr;
    float* yout = a.y_out ? reinterpret_cast<float*>(a.y_out) + row * H : nullptr;
    float ss = 0.0f;
    for (int i = lane; i < H; i += 32) {
        const float v = res ? __fadd_rn(res[i], x[i]) : x[i];
        if (hout) hout[i] = v;
        ss = __fmaf_rn(v, v, ss);
    }
    ss = warp_sum_f(ss);
    const float r = rsqrtf(__fadd_rn(__fdiv_rn(ss, (float)H), a.eps));
    float amax = 0.0f;
    for (int i = lane; i < H; i += 32) {
        const float v = res ? __fadd_rn(res[i], x[i]) : x[i];
        const float y = __fmul_rn(__fmul_rn(v, r), w[i]);          // x_row[i] * rrms * weight[i]   I64_rmsnorm.cu:147
        if (yout) yout[i] = y;
        amax = fmaxf(amax, fabsf(y));
    }
    if (a.q_out == nullptr) return;
    amax = warp_max(amax);
    const float scale = quant_scale(amax);
    for (int i = lane; i < H; i += 32) {
        const float v = res ? __fadd_rn(res[i], x[i]) : x[i];
        a.q_out[row * H + i] = (int8_t)quant_q(__fmul_rn(__fmul_rn(v, r), w[i]), scale);
    }
    if (lane == 0) a.scale_out[row] = scale;
}

}  // namespace

extern "C" int i64b_add_rmsnorm(const void* x, const void* residual, const void* weight, int dtype, float eps,
                                void* hidden_out, void* normed_out, int8_t* q_out, float* scale_out, int64_t n,
                                int hidden, i64b_stream_t stream) {
    if (n == 0) return I64B_OK;
    I64B_CHECK(x && weight, I64B_ERR_ARG, "null pointer");
    I64B_CHECK(dtype == I64B_BF16 || dtype == I64B_F32, I64B_ERR_ARG, "dtype must be BF16 or F32");
    I64B_CHECK(normed_out != nullptr || q_out != nullptr, I64B_ERR_ARG, "nothing to produce: pass normed_out and/or q_out");
    I64B_CHECK(q_out == nullptr || scale_out != nullptr, I64B_ERR_ARG, "q_out needs scale_out");
    I64B_CHECK(hidden_out == nullptr || residual != nullptr, I64B_ERR_ARG, "hidden_out is residual + x: pass the residual");
    I64B_CHECK(hidden > 0 && hidden % 8 == 0, I64B_ERR_ARG, "hidden size %d must be a multiple of 8", hidden);
    I64B_CHECK(dtype == I64B_F32 || hidden <= 32 * 8 * kMaxVec, I64B_ERR_ARG, "bf16 rows of up to %d elements", 32 * 8 * kMaxVec);
    NormArgs a = {x, residual, weight, hidden_out, normed_out, q_out, scale_out, n, hidden, eps};
    const int grid = (int)ceil_div64(n, kWarps);
    if (dtype == I64B_BF16)
        I64B_CUDA(i64b_launch(add_rmsnorm_bf16_kernel, dim3(grid), dim3(kWarps * 32), 0, (cudaStream_t)stream, a));
    else
        I64B_CUDA(i64b_launch(add_rmsnorm_f32_kernel, dim3(grid), dim3(kWarps * 32), 0, (cudaStream_t)stream, a));
    i64b_count_launch(1);
    return I64B_OK;
}
