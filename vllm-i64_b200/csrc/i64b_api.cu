// C ABI of the block: layer object (packed weights + tensor maps), forward orchestration
// (plan → fused scatter → grouped GEMM1+SwiGLU → [re-quant] → grouped GEMM2 → combine), stand-alone GEMM ops.
// No allocation, no synchronisation, no default-stream use; see include/i64b.h for the contract.
#include <stdarg.h>
#include <new>
#include <stdlib.h>

#include "i64b_fused.cuh"
#include "i64b_gemm.cuh"
#include "i64b_plan.cuh"

// ------------------------------------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void i64b_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void i64b_count_launch(int n) { g_launches += n; }
bool i64b_pdl_enabled() {
    static const bool on = !(getenv("I64B_PDL") && atoi(getenv("I64B_PDL")) == 0);
    return on;
}

static const int16_t kSiluLutQ7[I64B_SILU_LUT_SIZE] = {
#include "i64b_silu_lut.inc"
};

static size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// ------------------------------------------------------------------------------------------------ layer
static const int kNMmaChoices[3] = {64, 128, 256};

struct i64b_layer {
    i64b_moe_config cfg;
    int elt;                 // bytes per weight / activation element fed to the tensor cores
    // packed weights (caller-owned device memory)
    uint8_t* packed = nullptr;
    uint8_t* w13 = nullptr;      // [(E*2I + 2S), H]   rows = out-features (gate rows then up rows per expert; shared last)
    uint8_t* w2r = nullptr;      // [E*H, I]
    uint8_t* w2s = nullptr;      // [H, S]
    float* w13_scale = nullptr;  // W8A8 per out-row scales, same row indexing as the tensors above
    float* w2r_scale = nullptr;
    float* w2s_scale = nullptr;
    int16_t* silu_lut = nullptr;
    int* sync = nullptr;         // fused-kernel counters (persistent, zero between launches)
    int32_t* counts = nullptr;   // [K, E] per-route expert histogram of THIS layer's last forward (not in the shared workspace:
    int32_t* offsets = nullptr;  // last_expert_counts() must not return another layer's numbers) and its exclusive scan
    const uint8_t* table = nullptr;
    bool loaded = false;
    int debug = 0;               // I64B_DEBUG_* bits (i64b_moe_set_debug)
    int cta_limit = 0;           // CTAs the GEMM launches of this layer may use (0 = every SM; i64b_moe_set_cta_limit)
    const i64b_layer* next = nullptr;   // layer whose weights the fused decode kernel prefetches into L2
    CUtensorMap tm_w13[3], tm_w2r[3], tm_w2s[3];   // per n_mma choice (box rows differ)
    // activation tensor maps, valid for (ws_base, n)
    void* cached_ws = nullptr;
    int64_t cached_n = -1;
    CUtensorMap tm_xs, tm_inter_r, tm_inter_s;
    const void* cached_x = nullptr;   // fused decode kernel: maps over the caller's x
    int64_t cached_x_n = -1;
    CUtensorMap tm_x;
    CUtensorMap tm_xg;          // the same x with a one-row box (tile::gather4 source of GEMM1's routed rows)
    const void* cached_xg = nullptr;   // tm_xg has its own validity: it is encoded only when the opt-in gather path is taken
    int64_t cached_xg_n = -1;
    // carve of the last forward (debug taps)
    struct Carve {
        I64bPlanWs plan;
        int32_t *routes, *dest_row, *scatter_idx, *src_tok;
        uint8_t *xs, *inter_r, *inter_s;
        float *xs_scale, *inter_scale_r, *inter_scale_s, *y;
        int32_t *q14_r, *q14_s, *amax_r, *amax_s;
        int amax_ld_r, amax_ld_s;
        I64bTriWs tri;          // token-major layout (fused down projection)
        float* cat_scale;       // W8A8 token-major layout: [3, N] scales of the re-quantised intermediates by pair-order row
        size_t bytes;
    } cv;
    CUtensorMap tm_cat;         // [N, 2I+S] token-major intermediates over inter_r | inter_s (valid with cached_ws)
    // second stream of the layer: the dispatch row copies run there, under the shared expert's GEMM1 (overlap_copy())
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
};

// The bf16 prefill-sized forward hides its one pure-HBM pass — the dispatch row copies — under the shared expert's GEMM1,
// which reads the caller's x and needs only the dispatch INDICES: indices on the main stream, copies on the layer's second
// stream, shared-expert GEMM1 tiles first, routed tiles after the join.  Same kernels, same results (tests/test_gpu_tri.py).
// MEASURED SLOWER on B200 (profiles/r2s_*: 64K tokens 3.24 -> 3.44 ms/step, 16K tokens 0.914 -> 0.947): two half-sized
// GEMM1 launches each pay their own ramp and tail, and the copy takes HBM bandwidth from a GEMM1 that already runs near the
// SM operand-ingest limit.  OFF unless I64B_OVERLAP=1; the stream and events are made at weight load only when it is on.
static bool overlap_enabled() {
    static const bool on = getenv("I64B_OVERLAP") && atoi(getenv("I64B_OVERLAP")) != 0;
    return on;
}
static int ensure_side_stream(i64b_layer* L) {
    if (L->side != nullptr || !overlap_enabled()) return I64B_OK;
    I64B_CUDA(cudaStreamCreateWithFlags(&L->side, cudaStreamNonBlocking));
    I64B_CUDA(cudaEventCreateWithFlags(&L->ev_fork, cudaEventDisableTiming));
    I64B_CUDA(cudaEventCreateWithFlags(&L->ev_join, cudaEventDisableTiming));
    return I64B_OK;
}

static int w13_rows(const i64b_moe_config& c) { return c.num_experts * 2 * c.expert_inter + 2 * c.shared_inter; }

static void carve_packed(i64b_layer* L, uint8_t* base, size_t* total) {
    const i64b_moe_config& c = L->cfg;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        uint8_t* p = base ? base + off : nullptr;
        off += align256(bytes > 0 ? bytes : 16);
        return p;
    };
    L->w13 = take((size_t)w13_rows(c) * c.hidden_size * L->elt);
    L->w2r = take((size_t)c.num_experts * c.hidden_size * c.expert_inter * L->elt);
    L->w2s = take((size_t)c.hidden_size * c.shared_inter * L->elt);
    L->w13_scale = reinterpret_cast<float*>(take(sizeof(float) * w13_rows(c)));
    L->w2r_scale = reinterpret_cast<float*>(take(sizeof(float) * c.num_experts * c.hidden_size));
    L->w2s_scale = reinterpret_cast<float*>(take(sizeof(float) * c.hidden_size));
    L->silu_lut = reinterpret_cast<int16_t*>(take(sizeof(int16_t) * I64B_SILU_LUT_SIZE));
    L->sync = reinterpret_cast<int*>(take(sizeof(int) * I64B_SYNC_INTS));
    L->counts = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * I64B_MAX_BINS));
    L->offsets = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * I64B_MAX_BINS));
    if (total) *total = off;
}

static i64b_layer::Carve carve_ws(const i64b_layer* L, void* base, int64_t n) {
    const i64b_moe_config& c = L->cfg;
    i64b_layer::Carve cv;
    uint8_t* p = reinterpret_cast<uint8_t*>(base);
    const int64_t nn = n > 0 ? n : 1;
    const int64_t R = nn * c.top_k, T = R + nn;
    const bool q = c.mode == I64B_MODE_W8A8;
    const bool q8 = q || c.mode == I64B_MODE_FP8;      // FP8 keeps its bf16 SwiGLU rows where W8A8 keeps the Q14 integers
    cv.plan = i64b_plan_ws_carve(p, n, c.top_k, c.num_experts);
    size_t off = cv.plan.bytes;
    auto take = [&](size_t bytes) {
        uint8_t* r = p ? p + off : nullptr;
        off += align256(bytes > 0 ? bytes : 16);
        return r;
    };
    cv.routes = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * R));
    cv.dest_row = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * R));
    cv.scatter_idx = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * R));
    cv.src_tok = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * R));
    cv.xs = take((size_t)T * c.hidden_size * L->elt);
    cv.xs_scale = reinterpret_cast<float*>(take(sizeof(float) * T));
    cv.inter_r = take((size_t)R * c.expert_inter * L->elt);
    cv.inter_s = take((size_t)nn * c.shared_inter * L->elt);
    cv.inter_scale_r = reinterpret_cast<float*>(take(sizeof(float) * R));
    cv.inter_scale_s = reinterpret_cast<float*>(take(sizeof(float) * nn));
    // one |max| partial per (GEMM1 tile, epilogue column half): at most ceil(I/32) for BN = 32 and 64, and
    // 2*ceil(I/128) for BN = 128, which exceeds ceil(I/32) only for I <= 32
    cv.amax_ld_r = (c.expert_inter + 31) / 32 + 1;
    cv.amax_ld_s = (c.shared_inter + 31) / 32 + 1;
    cv.q14_r = reinterpret_cast<int32_t*>(take(q8 ? sizeof(int32_t) * R * c.expert_inter : 0));
    cv.q14_s = reinterpret_cast<int32_t*>(take(q8 ? sizeof(int32_t) * nn * c.shared_inter : 0));
    cv.amax_r = reinterpret_cast<int32_t*>(take(q ? sizeof(int32_t) * R * cv.amax_ld_r : 0));
    cv.amax_s = reinterpret_cast<int32_t*>(take(q ? sizeof(int32_t) * nn * cv.amax_ld_s : 0));
    const int64_t slabs = c.shared_inter > 0 ? (c.shared_inter + 511) / 512 : 1;
    cv.y = reinterpret_cast<float*>(take(sizeof(float) * (R + slabs * nn) * c.hidden_size));
    {
        const bool tri_shape = c.top_k == 2 && c.num_experts >= 1 && c.num_experts * c.num_experts <= I64B_MAX_EXPERTS;
        cv.tri = i64b_tri_ws_carve(p ? p + off : nullptr, tri_shape ? n : 0, c.top_k, tri_shape ? c.num_experts : 1);
        off += align256(cv.tri.bytes);
        cv.cat_scale = reinterpret_cast<float*>(take(q8 && tri_shape ? sizeof(float) * 3 * nn : 0));
    }
    cv.bytes = off;
    return cv;
}

// ------------------------------------------------------------------------------------------------ repack kernels
namespace {
// dst[(b*cols + c) * dst_ld_rows ... ] — batched transpose of 2-byte elements: src [B, R, C] -> dst rows
// (dst_row0 + b*dst_batch_rows + c), column r, pitch R.
__global__ void transpose16_kernel(const uint16_t* __restrict__ src, uint16_t* __restrict__ dst, int R, int C,
                                   int64_t dst_row0, int64_t dst_batch_rows) {
    __shared__ uint16_t tile[32][33];
    const int b = blockIdx.z;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const uint16_t* s = src + (int64_t)b * R * C;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        if (r < R && c < C) tile[i][threadIdx.x] = s[(int64_t)r * C + c];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (r < R && c < C) dst[(dst_row0 + (int64_t)b * dst_batch_rows + c) * R + r] = tile[threadIdx.x][i];
    }
}
}  // namespace

static int encode_weight_maps(i64b_layer* L) {
    const i64b_moe_config& c = L->cfg;
    const bool q = c.mode != I64B_MODE_BF16;        // one-byte operands (int8 or E4M3)
    for (int i = 0; i < 3; ++i) {
        const int n_mma = kNMmaChoices[i];
        int rc = i64b_encode_tmap(&L->tm_w13[i], L->w13, q, w13_rows(c), c.hidden_size,
                                  (int64_t)c.hidden_size * L->elt, n_mma / 2);
        if (rc) return rc;
        if (c.num_experts > 0) {
            rc = i64b_encode_tmap(&L->tm_w2r[i], L->w2r, q, (int64_t)c.num_experts * c.hidden_size, c.expert_inter,
                                  (int64_t)c.expert_inter * L->elt, n_mma);
            if (rc) return rc;
        } else {
            L->tm_w2r[i] = L->tm_w13[i];
        }
        if (c.shared_inter > 0) {
            rc = i64b_encode_tmap(&L->tm_w2s[i], L->w2s, q, c.hidden_size, c.shared_inter,
                                  (int64_t)c.shared_inter * L->elt, n_mma);
            if (rc) return rc;
        } else {
            L->tm_w2s[i] = L->tm_w2r[i];
        }
    }
    return I64B_OK;
}

static int pick_nmma(int64_t m_tiles_r, int n_r, int64_t m_tiles_s, int n_s, bool dual, const char* env) {
    const int forced = env_int(env, 0);
    if (forced == 64 || forced == 128 || forced == 256) return forced;
    const int sms = i64b_num_sms();
    for (int i = 2; i >= 0; --i) {
        const int bn = dual ? kNMmaChoices[i] / 2 : kNMmaChoices[i];
        const int64_t tiles = m_tiles_r * ((n_r + bn - 1) / bn) + m_tiles_s * ((n_s + bn - 1) / bn);
        if (2 * tiles >= sms) return kNMmaChoices[i];   // wider tiles re-read fewer activation bytes per weight byte
    }
    return 64;
}
static int nmma_index(int n_mma) { return n_mma == 64 ? 0 : (n_mma == 128 ? 1 : 2); }
// CTA pairs (cta_group::2) for prefill-sized problems: the 256-wide tile was picked and the groups are tall enough
// that padding every group to 256 rows wastes little.  I64B_CTA2=0 disables, =1 forces (when n_mma == 256).
static bool use_cta_pairs(int n_mma, int64_t rows_total, int groups) {
    static const int mode = env_int("I64B_CTA2", -1);
    if (n_mma != 256 || mode == 0) return false;
    if (mode == 1) return true;
    return rows_total >= (int64_t)groups * 1024;
}

extern "C" {

const char* i64b_last_error(void) { return g_err; }
int i64b_version(void) { return 100; }
int64_t i64b_launch_count(void) { return g_launches; }
void i64b_launch_count_reset(void) { g_launches = 0; }
const int16_t* i64b_silu_lut_q7(void) { return kSiluLutQ7; }

int i64b_moe_create(const i64b_moe_config* cfg, i64b_layer** out) {
    I64B_CHECK(cfg && out, I64B_ERR_ARG, "null argument");
    const i64b_moe_config& c = *cfg;
    I64B_CHECK(c.mode == I64B_MODE_BF16 || c.mode == I64B_MODE_W8A8 || c.mode == I64B_MODE_FP8, I64B_ERR_ARG, "unknown mode %d",
               c.mode);
    const int kalign = c.mode != I64B_MODE_BF16 ? 16 : 8;   // 16-byte TMA pitch
    I64B_CHECK(c.hidden_size > 0 && c.hidden_size % 16 == 0, I64B_ERR_ARG, "hidden_size %d must be a multiple of 16",
               c.hidden_size);
    I64B_CHECK(c.num_experts >= 0 && c.num_experts <= I64B_MAX_EXPERTS, I64B_ERR_ARG, "num_experts %d not in [0,%d]",
               c.num_experts, I64B_MAX_EXPERTS);
    I64B_CHECK(c.num_experts == 0 || (c.top_k >= 1 && c.top_k <= I64B_MAX_TOPK && c.top_k <= c.num_experts),
               I64B_ERR_ARG, "top_k (%d) cannot exceed num_experts (%d) or %d", c.top_k, c.num_experts, I64B_MAX_TOPK);
    I64B_CHECK(c.num_experts == 0 || (c.expert_inter > 0 && c.expert_inter % 16 == 0), I64B_ERR_ARG,
               "expert_inter %d must be a positive multiple of 16", c.expert_inter);
    I64B_CHECK(c.shared_inter >= 0 && c.shared_inter % 16 == 0, I64B_ERR_ARG, "shared_inter %d must be a multiple of 16",
               c.shared_inter);
    I64B_CHECK(c.num_experts > 0 || c.shared_inter > 0, I64B_ERR_ARG, "layer has neither experts nor a dense FFN");
    I64B_CHECK(c.num_experts == 0 || c.vocab_size > 0, I64B_ERR_ARG, "vocab_size must be positive");
    (void)kalign;
    i64b_layer* L = new (std::nothrow) i64b_layer();
    I64B_CHECK(L != nullptr, I64B_ERR_ARG, "out of host memory");
    L->cfg = c;
    if (c.num_experts == 0) L->cfg.top_k = 0;
    L->elt = c.mode != I64B_MODE_BF16 ? 1 : 2;
    *out = L;
    return I64B_OK;
}

void i64b_moe_destroy(i64b_layer* layer) {
    if (layer == nullptr) return;
    if (layer->ev_fork) cudaEventDestroy(layer->ev_fork);
    if (layer->ev_join) cudaEventDestroy(layer->ev_join);
    if (layer->side) cudaStreamDestroy(layer->side);
    delete layer;
}

int i64b_moe_set_scales(i64b_layer* L, const float* route_weight, int top_k, float routed_scale, float shared_scale) {
    I64B_CHECK(L && route_weight && top_k == L->cfg.top_k, I64B_ERR_ARG, "bad set_scales arguments");
    for (int k = 0; k < top_k; ++k) L->cfg.route_weight[k] = route_weight[k];
    L->cfg.routed_scale = routed_scale;
    L->cfg.shared_scale = shared_scale;
    return I64B_OK;
}

size_t i64b_moe_packed_bytes(const i64b_layer* layer) {
    if (!layer) return 0;
    i64b_layer tmp = *layer;
    size_t total = 0;
    carve_packed(&tmp, nullptr, &total);
    return total + 256;
}

int i64b_moe_load_bf16(i64b_layer* L, const void* gate, const void* up, const void* down, const void* shared_gate,
                       const void* shared_up, const void* shared_down, void* packed, i64b_stream_t stream) {
    I64B_CHECK(L && packed, I64B_ERR_ARG, "null argument");
    I64B_CHECK(L->cfg.mode == I64B_MODE_BF16, I64B_ERR_STATE, "layer was created in W8A8 mode");
    const i64b_moe_config& c = L->cfg;
    I64B_CHECK(c.num_experts == 0 || (gate && up && down), I64B_ERR_ARG, "expert weights missing");
    I64B_CHECK(c.shared_inter == 0 || (shared_gate && shared_up && shared_down), I64B_ERR_ARG, "shared weights missing");
    cudaStream_t s = (cudaStream_t)stream;
    L->packed = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(packed) + 255) & ~uintptr_t(255));
    carve_packed(L, L->packed, nullptr);
    const int H = c.hidden_size, I = c.expert_inter, S = c.shared_inter, E = c.num_experts;
    if (E > 0) {
        dim3 blk(32, 8);
        // gate/up [E,H,I] -> rows e*2I (+I) + i, pitch H
        dim3 g1((I + 31) / 32, (H + 31) / 32, E);
        transpose16_kernel<<<g1, blk, 0, s>>>(reinterpret_cast<const uint16_t*>(gate),
                                              reinterpret_cast<uint16_t*>(L->w13), H, I, 0, 2 * I);
        transpose16_kernel<<<g1, blk, 0, s>>>(reinterpret_cast<const uint16_t*>(up),
                                              reinterpret_cast<uint16_t*>(L->w13), H, I, I, 2 * I);
        // down [E,I,H] -> rows e*H + h, pitch I
        dim3 g2((H + 31) / 32, (I + 31) / 32, E);
        transpose16_kernel<<<g2, blk, 0, s>>>(reinterpret_cast<const uint16_t*>(down),
                                              reinterpret_cast<uint16_t*>(L->w2r), I, H, 0, H);
        I64B_LAUNCH_CHECK();
    }
    if (S > 0) {   // nn.Linear weights are already (out, in) = K-major
        uint8_t* sh = L->w13 + (size_t)E * 2 * I * H * 2;
        I64B_CUDA(cudaMemcpyAsync(sh, shared_gate, (size_t)S * H * 2, cudaMemcpyDeviceToDevice, s));
        I64B_CUDA(cudaMemcpyAsync(sh + (size_t)S * H * 2, shared_up, (size_t)S * H * 2, cudaMemcpyDeviceToDevice, s));
        I64B_CUDA(cudaMemcpyAsync(L->w2s, shared_down, (size_t)H * S * 2, cudaMemcpyDeviceToDevice, s));
    }
    I64B_CUDA(cudaMemsetAsync(L->sync, 0, sizeof(int) * I64B_SYNC_INTS, s));
    int rc = encode_weight_maps(L);
    if (rc) return rc;
    rc = ensure_side_stream(L);
    if (rc) return rc;
    L->loaded = true;
    L->cached_ws = nullptr;
    return I64B_OK;
}

static int load_bytes(i64b_layer* L, int mode, const int8_t* w13_q, const float* w13_s, const int8_t* w2_q, const float* w2_s,
                      const int8_t* sh13_q, const float* sh13_s, const int8_t* sh2_q, const float* sh2_s, void* packed,
                      i64b_stream_t stream) {
    I64B_CHECK(L && packed, I64B_ERR_ARG, "null argument");
    I64B_CHECK(L->cfg.mode == mode, I64B_ERR_STATE, "layer was created in mode %d, these weights are mode %d", L->cfg.mode, mode);
    const i64b_moe_config& c = L->cfg;
    I64B_CHECK(c.num_experts == 0 || (w13_q && w13_s && w2_q && w2_s), I64B_ERR_ARG, "expert weights missing");
    I64B_CHECK(c.shared_inter == 0 || (sh13_q && sh13_s && sh2_q && sh2_s), I64B_ERR_ARG, "shared weights missing");
    // the SwiGLU epilogue folds silu_integer's two out-of-range selects into `lut[clamp(g)] + max(g - 1024, 0)`
    I64B_CHECK(kSiluLutQ7[0] == 0 && kSiluLutQ7[I64B_SILU_LUT_SIZE - 1] == 1024, I64B_ERR_STATE,
               "Q7 SiLU table ends must be 0 and 1024");
    cudaStream_t s = (cudaStream_t)stream;
    L->packed = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(packed) + 255) & ~uintptr_t(255));
    carve_packed(L, L->packed, nullptr);
    const size_t H = c.hidden_size, I = c.expert_inter, S = c.shared_inter, E = c.num_experts;
    const cudaMemcpyKind dd = cudaMemcpyDeviceToDevice;
    if (E > 0) {
        I64B_CUDA(cudaMemcpyAsync(L->w13, w13_q, E * 2 * I * H, dd, s));
        I64B_CUDA(cudaMemcpyAsync(L->w13_scale, w13_s, sizeof(float) * E * 2 * I, dd, s));
        I64B_CUDA(cudaMemcpyAsync(L->w2r, w2_q, E * H * I, dd, s));
        I64B_CUDA(cudaMemcpyAsync(L->w2r_scale, w2_s, sizeof(float) * E * H, dd, s));
    }
    if (S > 0) {
        I64B_CUDA(cudaMemcpyAsync(L->w13 + E * 2 * I * H, sh13_q, 2 * S * H, dd, s));
        I64B_CUDA(cudaMemcpyAsync(L->w13_scale + E * 2 * I, sh13_s, sizeof(float) * 2 * S, dd, s));
        I64B_CUDA(cudaMemcpyAsync(L->w2s, sh2_q, H * S, dd, s));
        I64B_CUDA(cudaMemcpyAsync(L->w2s_scale, sh2_s, sizeof(float) * H, dd, s));
    }
    I64B_CUDA(cudaMemcpyAsync(L->silu_lut, kSiluLutQ7, sizeof(kSiluLutQ7), cudaMemcpyHostToDevice, s));
    I64B_CUDA(cudaMemsetAsync(L->sync, 0, sizeof(int) * I64B_SYNC_INTS, s));
    int rc = encode_weight_maps(L);
    if (rc) return rc;
    L->loaded = true;
    L->cached_ws = nullptr;
    return I64B_OK;
}

int i64b_moe_load_w8a8(i64b_layer* L, const int8_t* w13_q, const float* w13_s, const int8_t* w2_q, const float* w2_s,
                       const int8_t* sh13_q, const float* sh13_s, const int8_t* sh2_q, const float* sh2_s, void* packed,
                       i64b_stream_t stream) {
    return load_bytes(L, I64B_MODE_W8A8, w13_q, w13_s, w2_q, w2_s, sh13_q, sh13_s, sh2_q, sh2_s, packed, stream);
}

int i64b_moe_load_fp8(i64b_layer* L, const uint8_t* w13_q, const float* w13_s, const uint8_t* w2_q, const float* w2_s,
                      const uint8_t* sh13_q, const float* sh13_s, const uint8_t* sh2_q, const float* sh2_s, void* packed,
                      i64b_stream_t stream) {
    return load_bytes(L, I64B_MODE_FP8, reinterpret_cast<const int8_t*>(w13_q), w13_s, reinterpret_cast<const int8_t*>(w2_q),
                      w2_s, reinterpret_cast<const int8_t*>(sh13_q), sh13_s, reinterpret_cast<const int8_t*>(sh2_q), sh2_s,
                      packed, stream);
}

int i64b_moe_set_debug(i64b_layer* L, int flags) {
    I64B_CHECK(L, I64B_ERR_ARG, "null layer");
    L->debug = flags;
    return I64B_OK;
}

int i64b_moe_set_cta_limit(i64b_layer* L, int max_ctas) {
    I64B_CHECK(L, I64B_ERR_ARG, "null layer");
    I64B_CHECK(max_ctas >= 0, I64B_ERR_ARG, "negative CTA limit");
    L->cta_limit = max_ctas;
    return I64B_OK;
}

int i64b_moe_set_next(i64b_layer* L, const i64b_layer* next) {
    I64B_CHECK(L, I64B_ERR_ARG, "null layer");
    I64B_CHECK(next == nullptr || next->loaded, I64B_ERR_STATE, "next layer has no weights yet");
    L->next = next;
    return I64B_OK;
}

int i64b_moe_set_table(i64b_layer* L, const uint8_t* table_u8) {
    I64B_CHECK(L, I64B_ERR_ARG, "null layer");
    L->table = table_u8;
    return I64B_OK;
}

size_t i64b_moe_workspace_bytes(const i64b_layer* L, int64_t max_tokens) {
    if (!L) return 0;
    return carve_ws(L, nullptr, max_tokens).bytes + 1024;
}

// Fused combine policy (I64B_FUSE_COMBINE: 0 never, 1 always, unset = auto).  Measured on B200 (N = 16 384): for a dense
// FFN the fused epilogue replaces a 22 us pass for free; with routed experts the shared tiles' pull of 2 fp32 rows per
// token is not hidden under the MMAs (133 us fused vs 88 + 42 us), so auto keeps the separate combine pass there.
static int fuse_combine_mode() {
    static const int mode = getenv("I64B_FUSE_COMBINE") ? atoi(getenv("I64B_FUSE_COMBINE")) : -1;
    return mode;
}

// does GEMM2's epilogue finish the block (no separate combine launch)?
static bool combine_is_fused(const i64b_layer* L, int E, int K, int S) {
    const int fmode = (L->debug & I64B_DEBUG_FUSE_COMBINE) ? 1 : fuse_combine_mode();
    return S > 0 && fmode != 0 && (E == 0 || (fmode == 1 && K <= 2));
}

// Token-major fused down projection (i64b_launch_gemm_tri): GEMM2 + combine in one kernel, no fp32 per-assignment rows.
// Two forms.  When every route weight, the routed scale and the shared scale are powers of two (the named models), the
// scales are applied to the SwiGLU rows in GEMM1's epilogue (exact) and the down projection is ONE accumulation over
// K = 2*I_e + S (`one_acc`); otherwise the three segments keep their own accumulators and the epilogue folds them in the
// reference's operation order (bit-identical to GEMM2 + combine_kernel).
static bool is_pow2(float v) {
    int e = 0;
    return v > 0.0f && v < 3.0e38f && frexpf(v, &e) == 0.5f && e > -100 && e < 100;
}
static bool tri_one_acc(const i64b_layer* L, const i64b_moe_config& c) {
    static const int mode = getenv("I64B_TRI_ONEACC") ? atoi(getenv("I64B_TRI_ONEACC")) : -1;
    if (mode == 0 || (L->debug & I64B_DEBUG_TRI_FOLD) || c.mode != I64B_MODE_BF16) return false;
    for (int k = 0; k < c.top_k; ++k)
        if (!is_pow2(c.route_weight[k])) return false;
    return is_pow2(c.routed_scale) && (c.shared_inter == 0 || is_pow2(c.shared_scale));
}
// I64B_TRI: 0 never, 1 whenever the shape allows, unset = when the pair groups are tall enough: every group (one per
// UNORDERED expert pair, E(E-1)/2 of them with distinct routes) is padded to a multiple of 256 rows, ~128 wasted rows per
// group, which the saved y traffic (12 KB per token) and the combine pass pay for from a few hundred rows per group on.
static bool tri_selected(const i64b_layer* L, const i64b_moe_config& c, int64_t n, bool plain_call) {
    static const int mode = getenv("I64B_TRI") ? atoi(getenv("I64B_TRI")) : -1;
    static const int min_rows = getenv("I64B_TRI_MIN_ROWS") ? atoi(getenv("I64B_TRI_MIN_ROWS")) : 0;
    const int E = c.num_experts;
    if (mode == 0 || !plain_call || c.top_k != 2 || E < 2 || E * E > I64B_MAX_EXPERTS) return false;
    if (c.expert_inter % 64 != 0 || c.shared_inter % 64 != 0 || c.hidden_size % 256 != 0) return false;
    if (c.mode == I64B_MODE_W8A8 || c.mode == I64B_MODE_FP8) {     // int8 rows of 128 elements per k-block, a shared expert, rows the re-quantiser holds in registers
        static const int w8 = getenv("I64B_TRI_W8A8") ? atoi(getenv("I64B_TRI_W8A8")) : 1;
        if (!w8 || c.shared_inter <= 0 || c.expert_inter % 128 != 0 || c.shared_inter % 128 != 0 || c.expert_inter > 1536 ||
            c.shared_inter > 1536)
            return false;
    }
    if (L->debug & (I64B_DEBUG_KEEP_TAPS | I64B_DEBUG_FUSE_COMBINE | I64B_DEBUG_NO_TRI)) return false;
    if (mode == 1 || (L->debug & I64B_DEBUG_TRI)) return true;
    const int64_t pairs = (int64_t)E * (E - 1) / 2;
    const int64_t rows = min_rows > 0 ? min_rows : (tri_one_acc(L, c) ? I64B_TRI_MIN_ROWS_ONEACC : I64B_TRI_MIN_ROWS_FOLD);
    return n >= pairs * rows;
}

static int64_t mt_r_pre(int64_t R, int E) { return E > 0 ? R / 128 + E : 0; }

static bool fused_enabled() {
    static const bool on = !(getenv("I64B_FUSED") && atoi(getenv("I64B_FUSED")) == 0);
    return on;
}

struct FwdOpts {
    const int32_t* given_routes = nullptr;  // int32 [K,N] expert ids supplied by the caller
    int top_k = -1;                         // override (given routes only)
    const float* route_weight = nullptr;
    float routed_scale = 1.0f, shared_scale = 1.0f;
    bool use_shared = true;
    bool override_scales = false;
    unsigned stage_mask = 0xFFu;            // I64B_STAGE_* bits (per-stage timing replays)
    const float* preq_scale = nullptr;      // x is int8 [N,H] quantised by the caller (x_preq), these are its row scales
};

struct CtaLimitScope {          // the layer's GEMM CTA limit applies to the launches of one forward call
    int prev;
    explicit CtaLimitScope(int limit) : prev(i64b_set_gemm_cta_limit(limit)) {}
    ~CtaLimitScope() { i64b_set_gemm_cta_limit(prev); }
};

static int moe_forward_impl(i64b_layer* L, const void* x, const int64_t* token_ids, const FwdOpts& o,
                            void* out, int out_dtype, int64_t n, void* workspace, size_t workspace_bytes,
                            i64b_stream_t stream) {
    I64B_CHECK(L, I64B_ERR_ARG, "null layer");
    CtaLimitScope cta_scope(L->cta_limit);
    const int32_t* given_routes = o.given_routes;
    I64B_CHECK(L->loaded, I64B_ERR_STATE, "weights not loaded (call i64b_moe_load_bf16 / i64b_moe_load_w8a8)");
    I64B_CHECK(out_dtype == I64B_BF16 || out_dtype == I64B_F32, I64B_ERR_ARG, "unsupported output dtype %d", out_dtype);
    I64B_CHECK(n >= 0, I64B_ERR_ARG, "negative token count");
    if (n == 0) return I64B_OK;
    I64B_CHECK(x && out && workspace, I64B_ERR_ARG, "null pointer");
    i64b_moe_config c = L->cfg;
    if (o.top_k >= 0) {
        I64B_CHECK(given_routes != nullptr && o.top_k >= 1 && o.top_k <= I64B_MAX_TOPK && c.num_experts > 0,
                   I64B_ERR_ARG, "top_k override %d needs caller-supplied routes", o.top_k);
        c.top_k = o.top_k;
    }
    if (!o.use_shared) c.shared_inter = 0;
    if (o.override_scales) {
        for (int k = 0; k < c.top_k; ++k) c.route_weight[k] = o.route_weight ? o.route_weight[k] : 1.0f;
        c.routed_scale = o.routed_scale;
        c.shared_scale = o.shared_scale;
    }
    const bool q = c.mode == I64B_MODE_W8A8;
    const bool f8 = c.mode == I64B_MODE_FP8;        // E4M3 activations and weights, fp32 accumulation (core/fp8.py)
    const bool q8 = q || f8;                        // one-byte operands with per-row / per-channel scales
    const int gtype = q ? I64B_TYPE_INT8 : (f8 ? I64B_TYPE_FP8 : I64B_TYPE_BF16);
    const bool preq = o.preq_scale != nullptr;
    I64B_CHECK(!preq || q, I64B_ERR_ARG, "pre-quantised activations need a W8A8 layer");
    const int E = c.num_experts, K = c.top_k, H = c.hidden_size, I = c.expert_inter, S = c.shared_inter;
    I64B_CHECK(E == 0 || given_routes != nullptr || token_ids == nullptr || L->table != nullptr, I64B_ERR_STATE,
               "routing table not set (call i64b_moe_set_table)");
    I64B_CHECK(n * (K + 1) < ((int64_t)1 << 31) / 8, I64B_ERR_ARG, "token count %lld too large", (long long)n);
    void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~uintptr_t(1023));
    i64b_layer view = *L;     // carve with the effective (possibly overridden) geometry
    view.cfg = c;
    i64b_layer::Carve cv = carve_ws(&view, base, n);
    cv.plan.counts = L->counts;          // per-layer storage (the workspace is shared by all layers of a stream)
    cv.plan.offsets = L->offsets;
    I64B_CHECK(workspace_bytes >= cv.bytes + 1024, I64B_ERR_WORKSPACE, "workspace too small: %zu < %zu", workspace_bytes,
               cv.bytes + 1024);
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t R = n * K, T = R + n;

    const int64_t cache_key = n * 64 + K * 2 + (S > 0 ? 1 : 0);
    if (L->cached_ws != base || L->cached_n != cache_key) {
        int rc = i64b_encode_tmap(&L->tm_xs, cv.xs, q8, T, H, (int64_t)H * L->elt, 128);
        if (rc) return rc;
        if (E > 0) {
            rc = i64b_encode_tmap(&L->tm_inter_r, cv.inter_r, q8, R, I, (int64_t)I * L->elt, 128);
            if (rc) return rc;
        } else {
            L->tm_inter_r = L->tm_xs;
        }
        if (S > 0) {
            rc = i64b_encode_tmap(&L->tm_inter_s, cv.inter_s, q8, n, S, (int64_t)S * L->elt, 128);
            if (rc) return rc;
        } else {
            L->tm_inter_s = L->tm_inter_r;
        }
        if (E > 0 && K == 2 && (2 * I + S) * L->elt % 16 == 0) {
            rc = i64b_encode_tmap(&L->tm_cat, cv.inter_r, q8, n, 2 * I + S, (int64_t)(2 * I + S) * L->elt, 128);
            if (rc) return rc;
        }
        L->cached_ws = base;
        L->cached_n = cache_key;
    }
    L->cv = cv;
    const bool tri = tri_selected(L, c, n, given_routes == nullptr && !o.override_scales);
    cv.tri.ticket = cv.plan.sync + 1;

    // decode-sized bf16 batches: the whole layer in ONE persistent launch (csrc/i64b_fused.cu)
    if (fused_enabled() && !q8 && !(L->debug & I64B_DEBUG_NO_FUSED) && given_routes == nullptr && o.stage_mask == 0xFFu &&
        i64b_fused_supported((int)n, H, I, S, E, K)) {
        I64bFusedParams fp = {};
        fp.x = x;
        fp.token_ids = token_ids;
        fp.table = L->table;
        fp.n = (int)n; fp.hidden = H; fp.num_experts = E; fp.expert_inter = I; fp.shared_inter = S; fp.top_k = K;
        fp.vocab = c.vocab_size; fp.out_dtype = out_dtype;
        for (int k = 0; k < I64B_MAX_TOPK; ++k) fp.w[k] = k < K ? c.route_weight[k] : 0.0f;
        fp.routed_scale = c.routed_scale; fp.shared_scale = c.shared_scale;
        fp.inter_r = cv.inter_r; fp.inter_s = cv.inter_s; fp.y = cv.y; fp.out = out;
        fp.sync = L->sync;
        static const int prefetch_mode = env_int("I64B_PREFETCH", 0);
        if (L->next != nullptr && L->next != L && prefetch_mode > 0) {
            fp.next_w = L->next->w13;                              // w13 | w2r | w2s are contiguous in the pack
            fp.next_w_bytes = (uint64_t)(reinterpret_cast<const uint8_t*>(L->next->w13_scale) - L->next->w13);
            fp.prefetch_mode = prefetch_mode;
        }
        fp.tap_routes = cv.routes; fp.tap_dest = cv.dest_row; fp.tap_scatter = cv.scatter_idx;
        fp.tap_counts = cv.plan.counts; fp.tap_offsets = cv.plan.offsets;
        // tensor map over the caller's x (128-row tiles for the shared expert's GEMM1)
        if (L->cached_x != x || L->cached_x_n != n) {
            int rc = i64b_encode_tmap(&L->tm_x, x, false, n, H, (int64_t)H * 2, 128);
            if (rc) return rc;
            L->cached_x = x;
            L->cached_x_n = n;
        }
        return i64b_launch_fused(L->tm_x, L->tm_w13[1], L->tm_inter_r, L->tm_inter_s, L->tm_w2r[1], L->tm_w2s[1], fp, s);
    }

    // 1. routing + deterministic dispatch plan + GEMM group tables (device-resident, no host sync)
    I64bPlanArgs pa = {};
    pa.mode = given_routes ? I64B_ROUTE_GIVEN : (token_ids ? I64B_ROUTE_TABLE : I64B_ROUTE_NOIDS);
    pa.token_ids = token_ids;
    pa.given = given_routes;
    pa.table = L->table;
    pa.n = n;
    pa.vocab = c.vocab_size;
    pa.top_k = K;
    pa.num_experts = E;
    pa.expert_inter = I;
    pa.shared_inter = S;
    pa.hidden = H;
    // bf16: the shared expert (and a dense FFN) multiplies the caller's x in place; W8A8 needs the quantised copy unless
    // the caller already brings int8 rows (x_preq)
    const bool shared_from_x = (!q8 || preq) && S > 0;
    pa.shared_from_x = shared_from_x ? 1 : 0;
    pa.ticket = L->sync + 3 * I64B_SYNC_STRIDE;     // free slot of the layer's persistent counter area
    if (shared_from_x && (L->cached_x != x || L->cached_x_n != n)) {
        int rcx = i64b_encode_tmap(&L->tm_x, x, preq, n, H, (int64_t)H * L->elt, 128);
        if (rcx) return rcx;
        L->cached_x = x;
        L->cached_x_n = n;
    }
    int rc = I64B_OK;
    if (o.stage_mask & I64B_STAGE_PLAN) {
        static const int fuse_tri_plan = env_int("I64B_TRI_PLAN_FUSED", 1);     // 0: the pair sort as its own launch (A/B)
        rc = i64b_run_plan(pa, cv.plan, s, (tri && fuse_tri_plan) ? &cv.tri : nullptr);
        if (rc) return rc;
        if (tri && !fuse_tri_plan) {
            rc = i64b_run_tri_plan(cv.plan, cv.tri, n, E, H, s);
            if (rc) return rc;
        }
    }

    // 2. fused (quantize-and-)scatter
    I64bDispatchArgs da = {};
    da.x = x;
    da.n = n;
    da.hidden = H;
    da.top_k = K;
    da.num_experts = E;
    da.quantize = f8 ? 2 : ((q && !preq) ? 1 : 0);
    da.preq_scale = o.preq_scale;
    da.shared_copy = (S > 0 && !shared_from_x) ? 1 : 0;
    da.xs = cv.xs;
    da.xs_scale = cv.xs_scale;
    da.routes_out = cv.routes;
    da.scatter_idx_out = cv.scatter_idx;
    da.dest_row_out = cv.dest_row;
    da.tri = tri ? 1 : 0;
    da.tw = cv.tri;
    // GEMM1 without a dispatched copy: the pair kernel's producer gathers the routed rows from x with tile::gather4.  Bit-
    // identical (tests/test_gpu_tri.py) and MEASURED MUCH SLOWER on B200 (profiles/r2t_*: 64K tokens 3.24 -> 4.58 ms/step):
    // 32 four-row requests per k-block feed the SM at well under the rate of one 128-row box, and GEMM1 is bound by exactly
    // that operand ingest.  Opt-in only (I64B_GATHER=1 or I64B_DEBUG_GATHER on the layer).
    static const int gather_mode = env_int("I64B_GATHER", 0);
    const bool gather = (gather_mode != 0 || (L->debug & I64B_DEBUG_GATHER)) && !q8 && E > 0 && S > 0 && shared_from_x &&
                        o.stage_mask == 0xFFu && use_cta_pairs(256, T, E + 1) && pick_nmma(mt_r_pre(R, E), I, (n + 127) / 128, S, true, "I64B_GEMM1_NMMA") == 256;
    if (gather) {
        da.phase = 1;
        da.src_tok_out = cv.src_tok;
        if (L->cached_xg != x || L->cached_xg_n != n) {     // not tied to cached_x: the fused decode branch sets that one too
            static const int gbox = env_int("I64B_GATHER_BOX", 1);
            int rcx = i64b_encode_tmap(&L->tm_xg, x, false, n, H, (int64_t)H * 2, gbox);
            if (rcx) return rcx;
            L->cached_xg = x;
            L->cached_xg_n = n;
        }
    }
    // prefill-sized bf16 blocks with a shared expert: the row copies leave the main stream (see overlap_enabled())
    static const int overlap_min = env_int("I64B_OVERLAP_MIN_TOKENS", 4096);
    const bool overlap = !gather && L->side != nullptr && !(L->debug & I64B_DEBUG_NO_OVERLAP) && !q8 && E > 0 && S > 0 && shared_from_x && o.stage_mask == 0xFFu && n >= overlap_min;
    if (overlap) {
        da.phase = 1;
        rc = i64b_run_dispatch(da, cv.plan, s);
        if (rc) return rc;
        I64B_CUDA(cudaEventRecord(L->ev_fork, s));
        I64B_CUDA(cudaStreamWaitEvent(L->side, L->ev_fork, 0));
        da.phase = 2;
        rc = i64b_run_dispatch(da, cv.plan, L->side);
        if (rc) return rc;
        I64B_CUDA(cudaEventRecord(L->ev_join, L->side));
    } else if (o.stage_mask & I64B_STAGE_DISPATCH) {
        rc = i64b_run_dispatch(da, cv.plan, s);
        if (rc) return rc;
    }

    bool combine_fused = false;
    const int64_t mt_r = E > 0 ? R / 128 + E : 0;     // upper bound on routed m-tiles
    const int64_t mt_s = S > 0 ? (n + 127) / 128 : 0;
    const int64_t pt_r = E > 0 ? R / 256 + E : 0;     // the same in 256-row units (CTA pairs)
    const int64_t pt_s = S > 0 ? (n + 255) / 256 : 0;

    // 3. GEMM1: gate|up projections + fused SwiGLU epilogue
    {
        const int n_mma = pick_nmma(mt_r, I, mt_s, S, true, "I64B_GEMM1_NMMA");
        const int bn = n_mma / 2;
        const int nh = bn >= 64 ? 2 : 1;       // |max| partials per GEMM1 tile row: one per epilogue column half
        I64bGemmParams gp = {};
        gp.groups = cv.plan.groups1;
        gp.num_groups = E + 1;
        gp.k[0] = gp.k[1] = H;
        gp.out[0] = q8 ? (void*)cv.q14_r : (void*)cv.inter_r;      // FP8: bf16 SwiGLU rows, re-quantised below
        gp.out[1] = q8 ? (void*)cv.q14_s : (void*)cv.inter_s;
        gp.ld_out[0] = I;
        gp.ld_out[1] = S;
        gp.row_scale[0] = cv.xs_scale;
        gp.row_scale[1] = preq ? o.preq_scale : cv.xs_scale;      // shared expert on the caller's int8 rows: its scales, by token
        gp.col_scale[0] = gp.col_scale[1] = L->w13_scale;
        gp.amax_part[0] = cv.amax_r;
        gp.amax_part[1] = cv.amax_s;
        gp.amax_ld[0] = cv.amax_ld_r;
        gp.amax_ld[1] = cv.amax_ld_s;
        gp.silu_lut = L->silu_lut;
        if (tri && !q8) {     // SwiGLU rows go straight to their token's row of the [N, 2I+S] buffer
            gp.out[0] = gp.out[1] = cv.inter_r;
            gp.ld_out[0] = gp.ld_out[1] = 2 * I + S;
            gp.rowmap[0] = cv.tri.rowmap;
            gp.rowmap[1] = cv.tri.pos;
            gp.tri_col[0] = 0;
            gp.tri_col[1] = I;
            gp.tri_col[2] = 2 * I;
            gp.tri_mul[0] = gp.tri_mul[1] = gp.tri_mul[2] = 1.0f;
            if (tri_one_acc(L, c)) {        // powers of two: the products are exact
                gp.tri_mul[0] = c.route_weight[0] * c.routed_scale;
                gp.tri_mul[1] = c.route_weight[1] * c.routed_scale;
                gp.tri_mul[2] = S > 0 ? c.shared_scale : 1.0f;
                gp.tri_scaled = (gp.tri_mul[0] != 1.0f || gp.tri_mul[1] != 1.0f || gp.tri_mul[2] != 1.0f) ? 1 : 0;
            }
        }
        const int wi = nmma_index(n_mma);
        const bool pairs = use_cta_pairs(n_mma, T, E + (S > 0 ? 1 : 0));
        if (gather) {
            gp.gather_idx = cv.src_tok;
            gp.gather_rows = (int32_t)R;
            gp.n_tokens = (int32_t)n;
        }
        const int64_t tiles = pairs ? pt_r * ((I + bn - 1) / bn) + pt_s * ((S + bn - 1) / bn)
                                    : mt_r * ((I + bn - 1) / bn) + mt_s * ((S + bn - 1) / bn);
        if (overlap) {      // the shared expert's tiles first (they read x), the routed ones once the row copies have landed
            const int64_t tiles_s = (pairs ? pt_s : mt_s) * ((S + bn - 1) / bn);
            I64bGemmParams gs = gp;
            gs.groups = cv.plan.groups1 + E;
            gs.num_groups = 1;
            rc = i64b_launch_gemm(q, true, n_mma, pairs ? 2 : 1, L->tm_xs, L->tm_x, L->tm_w13[wi], L->tm_w13[wi], gs, tiles_s, s);
            if (rc) return rc;
            I64B_CUDA(cudaStreamWaitEvent(s, L->ev_join, 0));
            gp.num_groups = E;
            rc = i64b_launch_gemm(q, true, n_mma, pairs ? 2 : 1, L->tm_xs, L->tm_x, L->tm_w13[wi], L->tm_w13[wi], gp,
                                  tiles - tiles_s, s);
            if (rc) return rc;
        } else if (o.stage_mask & I64B_STAGE_GEMM1) {
            rc = i64b_launch_gemm(gtype, true, n_mma, pairs ? 2 : 1, gather ? L->tm_xg : L->tm_xs, shared_from_x ? L->tm_x : L->tm_xs,
                                  L->tm_w13[wi], L->tm_w13[wi], gp, tiles, s);
            if (rc) return rc;
        }
        if (f8 && (o.stage_mask & I64B_STAGE_REQUANT)) {  // quantize_activations_fp8 of the SwiGLU rows (fp8_linear, core/fp8.py:166)
            rc = i64b_run_requant_fp8(cv.q14_r, cv.inter_r, tri ? cv.cat_scale : cv.inter_scale_r, E > 0 ? R : 0, I, cv.q14_s,
                                      cv.inter_s, cv.inter_scale_s, S > 0 ? n : 0, S, s, tri ? &cv.tri : nullptr);
            if (rc) return rc;
        }
        if (q && (o.stage_mask & I64B_STAGE_REQUANT)) {   // re-quantise the integer SwiGLU output per row (quantize_activations_int8 of `inter`)
            rc = i64b_run_requant_q14(cv.q14_r, cv.amax_r, cv.amax_ld_r, ((I + bn - 1) / bn) * nh,
                                      reinterpret_cast<int8_t*>(cv.inter_r), tri ? cv.cat_scale : cv.inter_scale_r, E > 0 ? R : 0, I,
                                      cv.q14_s, cv.amax_s, cv.amax_ld_s, ((S + bn - 1) / bn) * nh,
                                      reinterpret_cast<int8_t*>(cv.inter_s), cv.inter_scale_s, S > 0 ? n : 0, S, s, nullptr, 0,
                                      tri ? &cv.tri : nullptr);
            if (rc) return rc;
        }
    }

    // 4'. token-major blocks: down projections of both routes and of the shared expert + the weighted combine, one kernel
    if (tri) {
        I64bGemmParams gp = {};
        gp.groups = cv.tri.groups;
        gp.num_groups = E * E;
        gp.k[0] = I;
        gp.k[1] = S;
        for (int k = 0; k < I64B_MAX_TOPK; ++k) gp.w[k] = k < K ? c.route_weight[k] : 0.0f;
        gp.routed_scale = c.routed_scale;
        gp.shared_scale = c.shared_scale;
        gp.final_out = out;
        gp.final_dtype = out_dtype;
        gp.ld_final = H;
        gp.row_tok = cv.tri.tok;
        gp.n_tokens = (int32_t)n;
        gp.row_scale[0] = cv.cat_scale;         // W8A8: [3, N] by pair-order row
        gp.col_scale[0] = L->w2r_scale;
        gp.col_scale[1] = L->w2s_scale;
        const int64_t units = n / 256 + (int64_t)E * E;
        if (o.stage_mask & I64B_STAGE_GEMM2) {
            rc = i64b_launch_gemm_tri(gtype, tri_one_acc(L, c), L->tm_cat, L->tm_w2r[1], L->tm_w2s[1], gp,
                                      units * ((H + 255) / 256), s);
            if (rc) return rc;
        }
        return I64B_OK;
    }

    // 4. GEMM2: down projections -> per-assignment fp32 outputs
    {
        const int n_mma = pick_nmma(mt_r + mt_s, H, 0, 0, false, "I64B_GEMM2_NMMA");
        I64bGemmParams gp = {};
        gp.groups = cv.plan.groups2;
        gp.num_groups = E + 1;
        gp.k[0] = E > 0 ? I : S;
        gp.k[1] = S > 0 ? S : I;
        gp.out[0] = gp.out[1] = cv.y;
        gp.ld_out[0] = gp.ld_out[1] = H;
        gp.out_dtype = I64B_F32;
        gp.epi = q8 ? I64B_EPI_DEQUANT : I64B_EPI_F32;
        gp.row_scale[0] = cv.inter_scale_r;
        gp.row_scale[1] = cv.inter_scale_s;
        gp.col_scale[0] = L->w2r_scale;
        gp.col_scale[1] = L->w2s_scale;
        // with a shared expert (or a dense FFN) its GEMM2 tiles finish the block: they pull the routed rows of
        // their tokens and write the combined output, so there is no separate combine pass
        combine_fused = combine_is_fused(L, E, K, S);
        if (combine_fused) {
            gp.fuse_combine = 1;
            gp.keep_shared_y = (L->debug & I64B_DEBUG_KEEP_TAPS) ? 1 : 0;
            gp.dest_row = cv.dest_row;
            gp.n_tokens = (int32_t)n;
            gp.top_k = E > 0 ? K : 0;
            for (int k = 0; k < I64B_MAX_TOPK; ++k) gp.w[k] = k < K ? c.route_weight[k] : 0.0f;
            gp.routed_scale = c.routed_scale;
            gp.shared_scale = c.shared_scale;
            gp.final_out = out;
            gp.final_dtype = out_dtype;
            gp.ld_final = H;
            gp.tile_done = cv.plan.sync;
        }
        const bool pairs = use_cta_pairs(n_mma, T, E + (S > 0 ? 1 : 0));
        const int wi = pairs ? 1 : nmma_index(n_mma);       // pairs: each CTA stages a 128-row half of the B tile
        const int64_t tiles = (pairs ? pt_r + pt_s : mt_r + mt_s) * ((H + n_mma - 1) / n_mma);
        if (o.stage_mask & I64B_STAGE_GEMM2) {
            rc = i64b_launch_gemm(gtype, false, n_mma, pairs ? 2 : 1, L->tm_inter_r, L->tm_inter_s, L->tm_w2r[wi],
                                  L->tm_w2s[wi], gp, tiles, s);
            if (rc) return rc;
        }
    }
    if (!(o.stage_mask & I64B_STAGE_COMBINE) || combine_fused) return I64B_OK;

    // 5. weighted un-permute combine with the shared / routed output scales
    I64bCombineArgs ca = {};
    ca.y = cv.y;
    ca.dest_row = cv.dest_row;
    ca.n = n;
    ca.hidden = H;
    ca.top_k = K;
    ca.has_shared = S > 0;
    ca.has_routed = E > 0;
    for (int k = 0; k < I64B_MAX_TOPK; ++k) ca.w[k] = k < K ? c.route_weight[k] : 0.0f;
    ca.routed_scale = c.routed_scale;
    ca.shared_scale = c.shared_scale;
    ca.out = out;
    ca.out_dtype = out_dtype;
    return i64b_run_combine(ca, s);
}

int i64b_moe_forward(i64b_layer* L, const void* x, const int64_t* token_ids, void* out, int out_dtype, int64_t n,
                     void* workspace, size_t workspace_bytes, i64b_stream_t stream) {
    return moe_forward_impl(L, x, token_ids, FwdOpts(), out, out_dtype, n, workspace, workspace_bytes, stream);
}

int i64b_moe_forward_preq(i64b_layer* L, const int8_t* x_q, const float* x_scale, const int64_t* token_ids, void* out,
                          int out_dtype, int64_t n, void* workspace, size_t workspace_bytes, i64b_stream_t stream) {
    I64B_CHECK(x_scale != nullptr || n == 0, I64B_ERR_ARG, "null scales");
    FwdOpts o;
    o.preq_scale = x_scale;
    return moe_forward_impl(L, x_q, token_ids, o, out, out_dtype, n, workspace, workspace_bytes, stream);
}

int i64b_fused_decode_supported(int n, int hidden, int expert_inter, int shared_inter, int num_experts, int top_k) {
    return i64b_fused_supported(n, hidden, expert_inter, shared_inter, num_experts, top_k);
}

int i64b_moe_uses_fused(const i64b_layer* L, int64_t n) {
    if (!L || !L->loaded || L->cfg.mode != I64B_MODE_BF16 || !fused_enabled() || (L->debug & I64B_DEBUG_NO_FUSED) || n > (1 << 20)) return 0;
    const i64b_moe_config& c = L->cfg;
    return i64b_fused_supported((int)n, c.hidden_size, c.expert_inter, c.shared_inter, c.num_experts, c.top_k);
}

int i64b_moe_fuses_combine(const i64b_layer* L, int64_t n) {
    if (!L || !L->loaded || i64b_moe_uses_fused(L, n)) return 0;
    if (tri_selected(L, L->cfg, n, true)) return 1;
    return combine_is_fused(L, L->cfg.num_experts, L->cfg.top_k, L->cfg.shared_inter) ? 1 : 0;
}

int i64b_moe_tri_form(const i64b_layer* L, int64_t n) {
    if (!L || !L->loaded || i64b_moe_uses_fused(L, n) || !tri_selected(L, L->cfg, n, true)) return 0;
    return tri_one_acc(L, L->cfg) ? 2 : 1;
}

int i64b_moe_forward_stages(i64b_layer* L, const void* x, const int64_t* token_ids, void* out, int out_dtype,
                            int64_t n, void* workspace, size_t workspace_bytes, unsigned stage_mask,
                            i64b_stream_t stream) {
    FwdOpts o;
    o.stage_mask = stage_mask;
    return moe_forward_impl(L, x, token_ids, o, out, out_dtype, n, workspace, workspace_bytes, stream);
}

int i64b_moe_forward_routed(i64b_layer* L, const void* x, const int32_t* routes, int top_k, const float* route_weight,
                            float routed_scale, float shared_scale, int use_shared, void* out, int out_dtype,
                            int64_t n, void* workspace, size_t workspace_bytes, i64b_stream_t stream) {
    I64B_CHECK(routes != nullptr || n == 0, I64B_ERR_ARG, "null routes");
    FwdOpts o;
    o.given_routes = routes;
    o.top_k = top_k;
    o.route_weight = route_weight;
    o.routed_scale = routed_scale;
    o.shared_scale = shared_scale;
    o.use_shared = use_shared != 0;
    o.override_scales = true;
    return moe_forward_impl(L, x, nullptr, o, out, out_dtype, n, workspace, workspace_bytes, stream);
}

size_t i64b_moe_workspace_bytes_routed(const i64b_layer* L, int64_t max_tokens, int top_k) {
    if (!L) return 0;
    i64b_layer view = *L;
    if (top_k > view.cfg.top_k) view.cfg.top_k = top_k;
    return carve_ws(&view, nullptr, max_tokens).bytes + 1024;
}

// ---- expert-parallel local compute (csrc/i64b_ep.cu holds the exchange)
int i64b_ep_build_groups(const i64b_ep_ctx* c, void* ep_workspace, int expert_inter, const I64bGroup** g1,
                         const I64bGroup** g2, int* num_groups, cudaStream_t stream);

struct DispatchedWs {
    uint8_t* inter;         // [rows, I] bf16 | int8
    int32_t* q14;           // W8A8: [rows, I] int32
    int32_t* amax;          // W8A8: [rows, amax_ld]
    float* inter_scale;     // W8A8: [rows]
    int amax_ld;
    size_t bytes;
};
static DispatchedWs carve_dispatched(const i64b_layer* L, const i64b_ep_ctx* c, void* workspace) {
    DispatchedWs w;
    uint8_t* base = workspace ? reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~uintptr_t(1023))
                              : nullptr;
    const size_t rows = (size_t)c->cap_src * c->size;
    const bool q = L->cfg.mode == I64B_MODE_W8A8;
    const int I = L->cfg.expert_inter;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        uint8_t* r = base ? base + off : nullptr;
        off += align256(bytes > 0 ? bytes : 16);
        return r;
    };
    w.inter = take(rows * I * L->elt);
    w.amax_ld = (I + 31) / 32 + 1;
    w.q14 = reinterpret_cast<int32_t*>(take(q ? rows * I * sizeof(int32_t) : 0));
    w.amax = reinterpret_cast<int32_t*>(take(q ? rows * w.amax_ld * sizeof(int32_t) : 0));
    w.inter_scale = reinterpret_cast<float*>(take(q ? rows * sizeof(float) : 0));
    w.bytes = off;
    return w;
}

size_t i64b_moe_dispatched_workspace_bytes(const i64b_layer* L, const i64b_ep_ctx* c) {
    if (!L || !c) return 0;
    return carve_dispatched(L, c, nullptr).bytes + 2048;
}

int i64b_moe_forward_dispatched(i64b_layer* L, const i64b_ep_ctx* c, void* ep_workspace, void* workspace,
                                size_t workspace_bytes, i64b_stream_t stream) {
    I64B_CHECK(L && c && ep_workspace && workspace, I64B_ERR_ARG, "null pointer");
    I64B_CHECK(L->loaded, I64B_ERR_STATE, "expert-parallel path: weights not loaded");
    CtaLimitScope cta_scope(L->cta_limit);
    const i64b_moe_config& cfg = L->cfg;
    I64B_CHECK(cfg.mode == c->x_mode, I64B_ERR_ARG, "layer mode %d does not match the transport mode %d", cfg.mode, c->x_mode);
    I64B_CHECK(cfg.num_experts * c->size == c->num_experts && cfg.hidden_size == c->hidden, I64B_ERR_ARG,
               "layer holds %d experts, EP context expects %d / %d", cfg.num_experts, c->num_experts, c->size);
    I64B_CHECK(c->size <= I64B_EP_MAX_RANKS, I64B_ERR_ARG, "at most %d EP ranks", I64B_EP_MAX_RANKS);
    I64B_CHECK(workspace_bytes >= i64b_moe_dispatched_workspace_bytes(L, c), I64B_ERR_WORKSPACE, "workspace too small");
    const bool q = cfg.mode == I64B_MODE_W8A8;
    const int H = cfg.hidden_size, I = cfg.expert_inter;
    const int64_t cap_rows = c->cap_src * c->size;
    const DispatchedWs w = carve_dispatched(L, c, workspace);
    void* xs = c->peer_xs[c->rank];
    const int64_t key = -(cap_rows + 7);
    if (L->cached_ws != xs || L->cached_n != key || L->cached_x != w.inter) {
        int rc = i64b_encode_tmap(&L->tm_xs, xs, q, cap_rows, H, (int64_t)H * L->elt, 128);
        if (rc) return rc;
        rc = i64b_encode_tmap(&L->tm_inter_r, w.inter, q, cap_rows, I, (int64_t)I * L->elt, 128);
        if (rc) return rc;
        L->cached_ws = xs;
        L->cached_n = key;
        L->cached_x = w.inter;
        L->cached_x_n = -1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const I64bGroup *g1 = nullptr, *g2 = nullptr;
    int ng = 0;
    int rc = i64b_ep_build_groups(c, ep_workspace, I, &g1, &g2, &ng, s);
    if (rc) return rc;
    // expected load (balanced routing): the real row counts live on the device only; one group per (expert, source)
    // (this rank's own assignment count stands in for the rows it receives; a rank without tokens still runs its experts)
    const int64_t rows_exp = (int64_t)c->n_local * c->top_k > 0 ? (int64_t)c->n_local * c->top_k : 1;
    const int64_t mt = rows_exp / 128 + (ng - 1), pt = rows_exp / 256 + (ng - 1);
    {
        const int n_mma = pick_nmma(mt, I, 0, 0, true, "I64B_GEMM1_NMMA");
        const int bn = n_mma / 2;
        const bool pairs = use_cta_pairs(n_mma, rows_exp, ng - 1);
        I64bGemmParams gp = {};
        gp.groups = g1;
        gp.num_groups = ng;
        gp.k[0] = gp.k[1] = H;
        gp.out[0] = gp.out[1] = q ? (void*)w.q14 : (void*)w.inter;
        gp.ld_out[0] = gp.ld_out[1] = I;
        gp.row_scale[0] = gp.row_scale[1] = c->peer_xs_scale[c->rank];
        gp.col_scale[0] = gp.col_scale[1] = L->w13_scale;
        gp.amax_part[0] = gp.amax_part[1] = w.amax;
        gp.amax_ld[0] = gp.amax_ld[1] = w.amax_ld;
        gp.silu_lut = L->silu_lut;
        const int wi = nmma_index(n_mma);
        // the grid is persistent: any upper bound >= the SM count gives the full machine
        rc = i64b_launch_gemm(q, true, n_mma, pairs ? 2 : 1, L->tm_xs, L->tm_xs, L->tm_w13[wi], L->tm_w13[wi], gp,
                              (pairs ? pt : mt) * ((I + bn - 1) / bn), s);
        if (rc) return rc;
        if (q) {
            const int nh = bn >= 64 ? 2 : 1;
            rc = i64b_run_requant_q14(w.q14, w.amax, w.amax_ld, ((I + bn - 1) / bn) * nh, reinterpret_cast<int8_t*>(w.inter),
                                      w.inter_scale, cap_rows, I, nullptr, nullptr, 0, 0, nullptr, nullptr, 0, 0, s, g1, ng - 1);
            if (rc) return rc;
        }
    }
    {
        const int n_mma = pick_nmma(mt, H, 0, 0, false, "I64B_GEMM2_NMMA");
        const bool pairs = use_cta_pairs(n_mma, rows_exp, ng - 1);
        I64bGemmParams gp = {};
        gp.groups = g2;
        gp.num_groups = ng;
        gp.k[0] = gp.k[1] = I;
        gp.out[0] = gp.out[1] = c->peer_y[c->rank];
        gp.ld_out[0] = gp.ld_out[1] = H;
        gp.out_dtype = c->y_dtype;
        gp.epi = q ? I64B_EPI_DEQUANT : I64B_EPI_F32;
        gp.row_scale[0] = gp.row_scale[1] = w.inter_scale;
        gp.col_scale[0] = gp.col_scale[1] = L->w2r_scale;
        const int wi = pairs ? 1 : nmma_index(n_mma);
        rc = i64b_launch_gemm(q, false, n_mma, pairs ? 2 : 1, L->tm_inter_r, L->tm_inter_r, L->tm_w2r[wi],
                              L->tm_w2r[wi], gp, (pairs ? pt : mt) * ((H + n_mma - 1) / n_mma), s);
        if (rc) return rc;
    }
    return I64B_OK;
}

int i64b_copy_device(void* dst, const void* src, size_t bytes, i64b_stream_t stream) {
    if (bytes == 0) return I64B_OK;
    I64B_CHECK(dst && src, I64B_ERR_ARG, "null pointer");
    I64B_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return I64B_OK;
}

int i64b_moe_forward_host(i64b_layer* L, const void* x_host, const int64_t* token_ids_host, void* out_host,
                          int out_dtype, int64_t n, void* x_dev, int64_t* ids_dev, void* out_dev, void* workspace,
                          size_t workspace_bytes, i64b_stream_t stream) {
    I64B_CHECK(L && x_host && out_host && x_dev && out_dev, I64B_ERR_ARG, "null pointer");
    I64B_CHECK(token_ids_host == nullptr || ids_dev != nullptr, I64B_ERR_ARG, "ids staging buffer missing");
    if (n == 0) return I64B_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t H = L->cfg.hidden_size;
    I64B_CUDA(cudaMemcpyAsync(x_dev, x_host, (size_t)n * H * 2, cudaMemcpyHostToDevice, s));
    if (token_ids_host)
        I64B_CUDA(cudaMemcpyAsync(ids_dev, token_ids_host, (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    int rc = i64b_moe_forward(L, x_dev, token_ids_host ? ids_dev : nullptr, out_dev, out_dtype, n, workspace,
                              workspace_bytes, stream);
    if (rc) return rc;
    I64B_CUDA(cudaMemcpyAsync(out_host, out_dev, (size_t)n * H * (out_dtype == I64B_F32 ? 4 : 2),
                              cudaMemcpyDeviceToHost, s));
    return I64B_OK;
}

void* i64b_moe_tap(i64b_layer* L, int what) {
    if (!L) return nullptr;
    const i64b_layer::Carve& cv = L->cv;
    switch (what) {
        case I64B_TAP_ROUTES: return cv.routes;
        case I64B_TAP_DEST_ROW: return cv.dest_row;
        case I64B_TAP_SCATTER_IDX: return cv.scatter_idx;
        case I64B_TAP_COUNTS: return cv.plan.counts;
        case I64B_TAP_OFFSETS: return cv.plan.offsets;
        case I64B_TAP_XS: return cv.xs;
        case I64B_TAP_XS_SCALE: return cv.xs_scale;
        case I64B_TAP_INTER_R: return cv.inter_r;
        case I64B_TAP_INTER_S: return cv.inter_s;
        case I64B_TAP_INTER_SCALE_R: return cv.inter_scale_r;
        case I64B_TAP_INTER_SCALE_S: return cv.inter_scale_s;
        case I64B_TAP_Q14_R: return cv.q14_r;
        case I64B_TAP_Q14_S: return cv.q14_s;
        case I64B_TAP_Y: return cv.y;
        case I64B_TAP_GROUP_ROWS: return cv.plan.group_start;
        default: return nullptr;
    }
}

// ------------------------------------------------------------------------------------------------ stand-alone GEMMs
static int pick_nmma_single(int64_t m, int64_t n) {
    const int forced = env_int("I64B_GEMM_NMMA", 0);
    if (forced == 64 || forced == 128 || forced == 256) return forced;
    const int sms = i64b_num_sms();
    const int64_t mt = (m + 127) / 128;
    for (int i = 2; i >= 0; --i)
        if (2 * mt * ((n + kNMmaChoices[i] - 1) / kNMmaChoices[i]) >= sms) return kNMmaChoices[i];
    return 64;
}

int i64b_int8_gemm(const int8_t* x_q, const float* x_scale, const int8_t* w_q, const float* w_scale, float* out,
                   int32_t* acc_out, int64_t m, int64_t n, int64_t k, i64b_stream_t stream) {
    if (m == 0 || n == 0) return I64B_OK;
    I64B_CHECK(x_q && w_q && (out || acc_out), I64B_ERR_ARG, "null pointer");
    I64B_CHECK(out == nullptr || (x_scale && w_scale), I64B_ERR_ARG, "scales required for the dequantised output");
    I64B_CHECK(k > 0 && k % 16 == 0 && n % 16 == 0, I64B_ERR_ARG, "k (%lld) and n (%lld) must be multiples of 16",
               (long long)k, (long long)n);
    I64B_CHECK(m < ((int64_t)1 << 31) && n < ((int64_t)1 << 31), I64B_ERR_ARG, "matrix too large");
    const int n_mma = pick_nmma_single(m, n);
    CUtensorMap ta, tb;
    int rc = i64b_encode_tmap(&ta, x_q, true, m, k, k, 128);
    if (rc) return rc;
    rc = i64b_encode_tmap(&tb, w_q, true, n, k, k, n_mma);
    if (rc) return rc;
    I64bGemmParams gp = {};
    gp.inline_group = I64bGroup{0, (int32_t)m, 0, (int32_t)n, 0, 0, 0, 0};
    gp.num_groups = 1;
    gp.k[0] = gp.k[1] = (int32_t)k;
    gp.out[0] = gp.out[1] = out;
    gp.ld_out[0] = gp.ld_out[1] = n;
    gp.out_dtype = I64B_F32;
    gp.epi = out ? I64B_EPI_DEQUANT : I64B_EPI_F32;
    gp.row_scale[0] = gp.row_scale[1] = x_scale;
    gp.col_scale[0] = gp.col_scale[1] = w_scale;
    gp.raw_out = acc_out;
    gp.ld_raw = n;
    const int64_t tiles = ((m + 127) / 128) * ((n + n_mma - 1) / n_mma);
    return i64b_launch_gemm(true, false, n_mma, 1, ta, ta, tb, tb, gp, tiles, (cudaStream_t)stream);
}

int i64b_fp8_gemm(const uint8_t* x_q, const float* x_scale, const uint8_t* w_q, const float* w_scale, void* out, int out_dtype,
                  int64_t m, int64_t n, int64_t k, i64b_stream_t stream) {
    if (m == 0 || n == 0) return I64B_OK;
    I64B_CHECK(x_q && w_q && out && x_scale && w_scale, I64B_ERR_ARG, "null pointer");
    I64B_CHECK(out_dtype == I64B_BF16 || out_dtype == I64B_F32, I64B_ERR_ARG, "unsupported output dtype");
    I64B_CHECK(k > 0 && k % 16 == 0 && n % 16 == 0, I64B_ERR_ARG, "k (%lld) and n (%lld) must be multiples of 16",
               (long long)k, (long long)n);
    I64B_CHECK(m < ((int64_t)1 << 31) && n < ((int64_t)1 << 31), I64B_ERR_ARG, "matrix too large");
    const int n_mma = pick_nmma_single(m, n);
    CUtensorMap ta, tb;
    int rc = i64b_encode_tmap(&ta, x_q, true, m, k, k, 128);
    if (rc) return rc;
    rc = i64b_encode_tmap(&tb, w_q, true, n, k, k, n_mma);
    if (rc) return rc;
    I64bGemmParams gp = {};
    gp.inline_group = I64bGroup{0, (int32_t)m, 0, (int32_t)n, 0, 0, 0, 0};
    gp.num_groups = 1;
    gp.k[0] = gp.k[1] = (int32_t)k;
    gp.out[0] = gp.out[1] = out;
    gp.ld_out[0] = gp.ld_out[1] = n;
    gp.out_dtype = out_dtype;
    gp.epi = I64B_EPI_DEQUANT;
    gp.row_scale[0] = gp.row_scale[1] = x_scale;
    gp.col_scale[0] = gp.col_scale[1] = w_scale;
    const int64_t tiles = ((m + 127) / 128) * ((n + n_mma - 1) / n_mma);
    return i64b_launch_gemm(I64B_TYPE_FP8, false, n_mma, 1, ta, ta, tb, tb, gp, tiles, (cudaStream_t)stream);
}

int i64b_bf16_gemm(const void* a, const void* b, const float* col_scale, void* out, int out_dtype, int64_t m,
                   int64_t n, int64_t k, i64b_stream_t stream) {
    if (m == 0 || n == 0) return I64B_OK;
    I64B_CHECK(a && b && out, I64B_ERR_ARG, "null pointer");
    I64B_CHECK(out_dtype == I64B_BF16 || out_dtype == I64B_F32, I64B_ERR_ARG, "unsupported output dtype");
    I64B_CHECK(k > 0 && k % 8 == 0 && n % 16 == 0, I64B_ERR_ARG, "k (%lld) must be a multiple of 8 and n (%lld) of 16",
               (long long)k, (long long)n);
    I64B_CHECK(m < ((int64_t)1 << 31) && n < ((int64_t)1 << 31), I64B_ERR_ARG, "matrix too large");
    const int n_mma = pick_nmma_single(m, n);
    CUtensorMap ta, tb;
    int rc = i64b_encode_tmap(&ta, a, false, m, k, k * 2, 128);
    if (rc) return rc;
    rc = i64b_encode_tmap(&tb, b, false, n, k, k * 2, n_mma);
    if (rc) return rc;
    I64bGemmParams gp = {};
    gp.inline_group = I64bGroup{0, (int32_t)m, 0, (int32_t)n, 0, 0, 0, 0};
    gp.num_groups = 1;
    gp.k[0] = gp.k[1] = (int32_t)k;
    gp.out[0] = gp.out[1] = out;
    gp.ld_out[0] = gp.ld_out[1] = n;
    gp.out_dtype = out_dtype;
    gp.epi = col_scale ? I64B_EPI_DEQUANT : I64B_EPI_F32;
    gp.col_scale[0] = gp.col_scale[1] = col_scale;
    const int64_t tiles = ((m + 127) / 128) * ((n + n_mma - 1) / n_mma);
    return i64b_launch_gemm(false, false, n_mma, 1, ta, ta, tb, tb, gp, tiles, (cudaStream_t)stream);
}

}  // extern "C"
