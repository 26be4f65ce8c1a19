"""f4 — the block in FP8 (E4M3 weights, dynamically quantised E4M3 activations, fp32 accumulation on tcgen05 kind::f8f6f4),
semantics of the reference's core/fp8.py + layers/dense_mlp.py:75-83 applied per expert.

Bars: the E4M3 activation bytes and their scales are BIT-EXACT against the oracle (and through tests/golden/tiny_fp8.npz
against the reference's own quantize_activations_fp8); block outputs agree with the oracle's FP8 block within a stated
tolerance (the fp32 summation order of the tensor core is not the oracle's, and an intermediate that lands on the other side
of a bf16 / E4M3 rounding boundary moves one term of a 512-term sum by a few per cent), and sit as close to the fp32 block as
the oracle's FP8 block does."""
import numpy as np
import pytest
import torch

import i64_test_helpers as Hh
from oracle import i64_oracle as O

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("CUDA device required", allow_module_level=True)

from vllm_i64 import _C
from vllm_i64.kernels.block import I64Block

TOL_MAX, TOL_MEAN = 3e-2, 3e-3        # vs the FP8 oracle, relative to max|ref| (E4M3 carries 3 mantissa bits: 6 % steps)


def build_fp8(cfg, wq, table):
    dev = torch.device("cuda", torch.cuda.current_device())
    blk = I64Block(cfg.hidden_size, cfg.num_experts, cfg.expert_inter, cfg.shared_inter, cfg.vocab_size, cfg.top_k,
                   cfg.route_weights(), cfg.routed_output_scale, cfg.shared_output_scale, _C.MODE_FP8, dev)
    enc = lambda a: None if a is None else Hh.to_cuda(O.e4m3_encode(a))
    f32 = lambda a: None if a is None else Hh.to_cuda(np.ascontiguousarray(a, dtype=np.float32))
    blk.load_fp8(enc(wq.w13_q), f32(wq.w13_s), enc(wq.w2_q), f32(wq.w2_s), enc(wq.sh13_q), f32(wq.sh13_s), enc(wq.sh2_q),
                 f32(wq.sh2_s))
    if cfg.num_experts:
        blk.set_table(torch.from_numpy(np.asarray(table, dtype=np.int64)))
    return blk


CASES = [
    dict(id="small", shape=(256, 4, 128, 128, 1000, 2), n=333),
    dict(id="noshared-zipf", shape=(256, 8, 128, 0, 5000, 2), n=1501, zipf=True),
    dict(id="top1", shape=(512, 4, 256, 256, 1000, 1), n=700),
    dict(id="config3-shape", shape=(1024, 8, 512, 1280, 32000, 2), n=4096, sample=96),
    dict(id="config3-16k", shape=(1024, 8, 512, 1280, 32000, 2), n=16384, sample=64),
]


@pytest.mark.parametrize("case", CASES, ids=[c["id"] for c in CASES])
def test_fp8_block_matches_oracle(case):
    h, e, ie, s, v, k = case["shape"]
    n = case["n"]
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=9, zipf=case.get("zipf", False))
    wq = O.quantize_moe_weights_fp8(w)
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    blk = build_fp8(cfg, wq, table)
    blk.set_debug(_C.DEBUG_KEEP_TAPS)
    out = blk.forward(xg, ig, torch.float32)
    again = blk.forward(xg, ig, torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(out, again), "not run-to-run identical"
    # the dispatched E4M3 rows and their scales: bit-exact
    xq, xs = O.quantize_rows_fp8(x)
    dest = blk.tap(_C.TAP_DEST_ROW, (k, n), torch.int32).cpu().numpy()
    got_q = blk.tap(_C.TAP_XS, (n * k + n, h), torch.uint8).cpu().numpy()
    got_s = blk.tap(_C.TAP_XS_SCALE, (n * k + n,), torch.float32).cpu().numpy()
    want_q = O.e4m3_encode(xq)
    for r in range(k):
        assert np.array_equal(got_q[dest[r]], want_q) and np.array_equal(got_s[dest[r]], xs[:, 0])
    if s:
        assert np.array_equal(got_q[n * k:], want_q) and np.array_equal(got_s[n * k:], xs[:, 0])
    sample = case.get("sample")
    rows = np.arange(n) if not sample else np.unique(np.linspace(0, n - 1, sample).astype(int))
    ref = O.moe_block_fp8(x[rows], ids[rows], table, wq, cfg)
    got = out.cpu().numpy()[rows]
    mx, mean = O.rel_errors(got, ref)
    assert mx <= TOL_MAX and mean <= TOL_MEAN, (mx, mean)
    # and as close to the un-quantised block as the oracle's FP8 block is
    f32 = O.moe_block_fp32(x[rows], ids[rows], table, w, cfg)
    e_ref, e_got = O.rel_errors(ref, f32), O.rel_errors(got, f32)
    assert e_got[1] <= 1.25 * e_ref[1] + 1e-4, (e_got, e_ref)
    print(f"fp8 {case['id']}: vs fp8 oracle max-rel {mx:.2e} mean-rel {mean:.2e}; vs fp32 {e_got[1]:.2e} (oracle fp8 {e_ref[1]:.2e})")


def test_fp8_dense_ffn_and_bf16_output():
    """num_experts == 0: the dense FFN of layers/dense_mlp.py:75-83 (dense-306 width), bf16 output, graph replay."""
    n = 513
    rng = np.random.RandomState(2)
    u = lambda shape, fan: Hh.bf16_round(rng.uniform(-1, 1, size=shape).astype(np.float32) / np.sqrt(fan))
    w = O.MoEWeights(np.zeros((0, 1024, 0), np.float32), np.zeros((0, 1024, 0), np.float32), np.zeros((0, 0, 1024), np.float32),
                     u((4096, 1024), 1024), u((4096, 1024), 1024), u((1024, 4096), 4096))
    cfg = O.MoEConfig(1024, 0, 0, 1000, top_k=0, shared_inter=4096)
    table = np.zeros((0, 1000), np.int64)
    x = Hh.bf16_round(rng.randn(n, 1024).astype(np.float32))
    wq = O.quantize_moe_weights_fp8(w)
    xg = Hh.to_cuda(x, torch.bfloat16)
    blk = build_fp8(cfg, wq, table)
    out = blk.forward(xg, None, torch.bfloat16)
    torch.cuda.synchronize()
    ref = O.moe_block_fp8(x, None, table, wq, cfg)
    mx, mean = O.rel_errors(out.float().cpu().numpy(), ref)
    assert mx <= TOL_MAX and mean <= 4e-3, (mx, mean)
    buf = torch.empty_like(out)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        blk.forward(xg, None, torch.bfloat16, out=buf)
    st.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        blk.forward(xg, None, torch.bfloat16, out=buf)
    buf.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(buf, out)


def test_fp8_layer_rejects_other_weights():
    cfg, w, table, ids, x = Hh.make_block(256, 4, 128, 128, 1000, 2, 8, seed=1)
    blk = build_fp8(cfg, O.quantize_moe_weights_fp8(w), table)
    with pytest.raises(RuntimeError):
        blk.load_bf16(*[Hh.to_cuda(t) for t in (w.gate, w.up, w.down, w.shared_gate, w.shared_up, w.shared_down)])


def test_fp8_host_mirror_ops():
    """core/fp8.py surface on the kernels: quantisers bit-exact with torch's own E4M3 conversion (= the reference's ops),
    fp8_linear / fp8_fused_gate_up against the restatement, TokenRoutedMLP.quantize_fp8() and DenseMLP.quantize_fp8()."""
    from vllm_i64.core import fp8 as F8
    from vllm_i64.layers.dense_mlp import DenseMLP
    from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP
    g = torch.Generator(device="cuda").manual_seed(3)
    x = (torch.randn(77, 256, device="cuda", generator=g) * torch.logspace(-2, 1, 77, device="cuda").unsqueeze(1)).to(torch.bfloat16)
    xq, xs = F8.quantize_activations_fp8(x)
    # the reference's expression evaluated by torch on the HOST (IEEE division; torch's CUDA kernels turn `/ 448.0` into a
    # multiplication by the rounded reciprocal, one ulp away on some rows — the golden vectors are the host's)
    xf = x.float().cpu()
    sc = xf.abs().amax(-1, keepdim=True).clamp(min=1e-12) / 448.0
    want = (xf / sc).clamp(-448, 448).to(torch.float8_e4m3fn)
    assert xq.dtype == torch.float8_e4m3fn and xs.shape == (77, 1)
    assert torch.equal(xs.cpu(), sc) and torch.equal(xq.view(torch.uint8).cpu(), want.view(torch.uint8))
    w = torch.randn(96, 256, device="cuda", generator=g) / 16
    wq, ws = F8.quantize_fp8(w)
    qo, so = O.quantize_rows_fp8(w.cpu().numpy())
    assert np.array_equal(wq.view(torch.uint8).cpu().numpy(), O.e4m3_encode(qo)) and np.array_equal(ws.cpu().numpy(), so[:, 0])
    y = F8.fp8_linear(x, wq, ws)
    assert y.dtype == x.dtype
    aq, as_ = O.quantize_rows_fp8(x.float().cpu().numpy())
    ref = O.scaled_mm(aq, as_, qo, so[:, 0])
    mx, mean = O.rel_errors(y.float().cpu().numpy(), ref)
    assert mx <= 1e-2 and mean <= 2e-3, (mx, mean)            # bf16 output rounding
    gate, up = F8.fp8_fused_gate_up(x.float(), wq, ws, 48)
    mx, mean = O.rel_errors(torch.cat([gate, up], -1).cpu().numpy(), ref)
    assert mx <= 1e-5 and mean <= 1e-6, (mx, mean)            # fp32 in, fp32 out: summation order only
    # modules
    cfg, wts, table, ids, xx = Hh.make_block(256, 4, 128, 128, 1000, 2, 300, seed=4)
    m = TokenRoutedMLP(256, 4 * 128, 4, 1000, shared_expert=True, shared_intermediate_size=128, top_k=2,
                       top_k_primary_weight=0.5, shared_output_scale=1.0, routed_output_scale=2.0)
    with torch.no_grad():
        m.gate_proj_w.copy_(torch.from_numpy(wts.gate)); m.up_proj_w.copy_(torch.from_numpy(wts.up))
        m.down_proj_w.copy_(torch.from_numpy(wts.down))
        m.shared_gate.weight.copy_(torch.from_numpy(wts.shared_gate)); m.shared_up.weight.copy_(torch.from_numpy(wts.shared_up))
        m.shared_down.weight.copy_(torch.from_numpy(wts.shared_down))
        m.topk_token_to_expert.copy_(torch.from_numpy(table)); m.token_to_expert.copy_(torch.from_numpy(table[0]))
    m = m.cuda().quantize_fp8()
    out = m(Hh.to_cuda(xx, torch.bfloat16).float(), Hh.to_cuda(ids))
    ref = O.moe_block_fp8(xx, ids, table, O.quantize_moe_weights_fp8(wts), cfg)
    mx, mean = O.rel_errors(out.float().cpu().numpy(), ref)
    assert mx <= TOL_MAX and mean <= TOL_MEAN, (mx, mean)
    d = DenseMLP(256, 512).cuda()
    xd = torch.randn(50, 256, device="cuda", generator=g)
    want = d(xd)
    got = d.quantize_fp8()(xd)
    mx, mean = O.rel_errors(got.cpu().numpy(), want.cpu().numpy())
    assert mean <= 0.1 and mx <= 0.35, (mx, mean)             # FP8 against the bf16 block: the format's own 6 % noise


@pytest.mark.parametrize("n,shape", [(900, (256, 4, 128, 128, 1000, 2)), (3001, (256, 8, 128, 256, 5000, 2)),
                                     (16384, (1024, 8, 512, 1280, 32000, 2))], ids=["small", "zipf", "config3-16k"])
def test_fp8_fused_down_projection_is_bit_identical_to_the_chain(n, shape):
    """The FP8 form of the token-major fused down projection (three fp32 accumulators, every segment dequantised with its own
    row and column scales, folded in the reference's order): the same fp32 operations as GEMM2 + combine_kernel."""
    h, e, ie, s, v, k = shape
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=13, zipf=(n == 3001))
    wq = O.quantize_moe_weights_fp8(w)
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    chain = build_fp8(cfg, wq, table)
    chain.set_debug(_C.DEBUG_NO_TRI)
    tri = build_fp8(cfg, wq, table)
    tri.set_debug(_C.DEBUG_TRI)
    assert chain.tri_form(n) == 0 and tri.tri_form(n) == 1
    for odt in (torch.float32, torch.bfloat16):
        want = chain.forward(xg, ig, odt)
        got = tri.forward(xg, ig, odt)
        torch.cuda.synchronize()
        bad = (got != want).any(dim=1).nonzero().flatten()
        assert bad.numel() == 0, f"{bad.numel()} of {n} rows differ from the FP8 chain, first {bad[:8].tolist()}"
    auto = build_fp8(cfg, wq, table)
    assert auto.tri_form(n) == (1 if n >= (e * (e - 1) // 2) * _C.TRI_MIN_ROWS_FOLD else 0)
