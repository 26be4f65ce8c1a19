"""Token-major fused down projection (i64b_launch_gemm_tri): GEMM2 of both routes and of the shared expert + the weighted
combine in ONE kernel, no fp32 per-assignment rows in memory.  Two forms:
* three accumulators folded in the epilogue: the same fp32 operations in the same order as GEMM2 + combine_kernel, so the
  bar is BIT IDENTITY with that chain (and the chain's own bar against the oracle);
* one accumulator (route weights x routed scale and the shared scale all powers of two — the named models): the scales go
  into the SwiGLU rows in GEMM1's epilogue (exact) and the block output is one sum over K = 2*I_e + S.  Same products, a
  different fp32 summation order: the bar is the oracle's, plus a few fp32 ulps from the chain.
Cases: small and config-5 geometry, no shared expert, skewed ids (empty / tiny / huge pair groups), both orders of a pair
and colliding routes (e0 == e1), token counts that are not a multiple of any tile, both output dtypes, unequal route
weights, scales other than 1, graph replay."""
import os

import numpy as np
import pytest
import torch

import i64_test_helpers as Hh
from oracle import i64_oracle as O

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("CUDA device required", allow_module_level=True)

from vllm_i64 import _C

CASES = [
    # h, e, ie, s, v, k, n, zipf, pw
    dict(id="small", shape=(256, 4, 64, 128, 1000, 2), n=700),
    dict(id="small-noshared", shape=(256, 4, 64, 0, 1000, 2), n=517),
    dict(id="tiny-n", shape=(256, 8, 64, 128, 1000, 2), n=5),
    dict(id="zipf", shape=(256, 8, 128, 192, 5000, 2), n=3001, zipf=True),
    dict(id="collisions", shape=(256, 4, 64, 128, 1000, 2), n=900, collide=3),
    dict(id="weights-95-05", shape=(256, 4, 64, 128, 1000, 2), n=1300, pw=0.95),
    dict(id="config5-8k", shape=(1024, 8, 512, 1280, 32000, 2), n=8192, sample=64),
    dict(id="config5-33k", shape=(1024, 8, 512, 1280, 32000, 2), n=33000, sample=32),
]
# the single-accumulator form needs power-of-two scales: (route weight, routed scale, shared scale)
ONEACC_SCALES = [(0.5, 2.0, 1.0), (0.5, 1.0, 0.5), (0.5, 4.0, 2.0)]


def _run(case, out_dtype, fold=True, scales=None):
    h, e, ie, s, v, k = case["shape"]
    n = case["n"]
    pw, rs, ss = scales if scales else (case.get("pw", 0.5), 2.0, 1.0)
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=11, zipf=case.get("zipf", False), pw=pw, rs=rs, ss=ss)
    if case.get("collide"):          # some vocabulary entries route both slots to the same expert
        table = np.array(table, copy=True)
        table[1, ::case["collide"]] = table[0, ::case["collide"]]
    routes = O.route_table(ids, table, n, e)
    if n > 100:
        assert (routes[0] < routes[1]).any() and (routes[0] > routes[1]).any(), "both orders of a pair must occur"
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    chain = Hh.build_gpu_block(cfg, w, table, "bf16")
    chain.set_debug(_C.DEBUG_NO_FUSED | _C.DEBUG_NO_TRI)
    tri = Hh.build_gpu_block(cfg, w, table, "bf16")
    tri.set_debug(_C.DEBUG_NO_FUSED | _C.DEBUG_TRI | (_C.DEBUG_TRI_FOLD if fold else 0))
    assert chain.tri_form(n) == 0
    assert tri.tri_form(n) == (1 if fold or pw != 0.5 else 2)
    want = chain.forward(xg, ig, out_dtype)
    got = tri.forward(xg, ig, out_dtype)
    again = tri.forward(xg, ig, out_dtype)
    torch.cuda.synchronize()
    assert torch.equal(got, again), "not run-to-run identical"
    if tri.tri_form(n) == 1:
        bad = (got != want).any(dim=1).nonzero().flatten()
        assert bad.numel() == 0, f"{bad.numel()} of {n} rows differ from the GEMM2 + combine chain, first {bad[:8].tolist()}"
    else:
        # same products, one fp32 accumulation instead of three rounded partial sums: a few fp32 ulps of the row's scale
        # (bf16 output: at most one bf16 ulp where the fp32 values straddle a rounding boundary)
        d = (got.float() - want.float()).abs().max().item()
        top = want.float().abs().max().item()
        assert d <= top * (2e-5 if out_dtype == torch.float32 else 2.0 ** -7), (d, top)
    sample = case.get("sample")
    rows = np.arange(n) if not sample else np.unique(np.linspace(0, n - 1, sample).astype(int))
    ref = O.moe_block_fp32(x[rows], ids[rows], table, w, cfg, inter_bf16=True)
    mx, mean = O.rel_errors(got.float().cpu().numpy()[rows], ref)
    bar_max, bar_mean = (Hh.TOL_MAX, Hh.TOL_MEAN) if out_dtype == torch.float32 else (2e-2, 4e-3)   # + one bf16 rounding of the output
    assert mx <= bar_max and mean <= bar_mean, (mx, mean)
    return tri, xg, ig, got


@pytest.mark.parametrize("case", CASES, ids=[c["id"] for c in CASES])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_tri_equals_chain(case, out_dtype):
    _run(case, out_dtype, fold=True)


@pytest.mark.parametrize("case", [c for c in CASES if "pw" not in c], ids=[c["id"] for c in CASES if "pw" not in c])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_tri_one_accumulator(case, out_dtype):
    _run(case, out_dtype, fold=False)


@pytest.mark.parametrize("scales", ONEACC_SCALES[1:], ids=["half", "x2"])
def test_tri_one_accumulator_scales_folded_into_gemm1(scales):
    _run(CASES[0], torch.float32, fold=False, scales=scales)
    _run(CASES[3], torch.bfloat16, fold=False, scales=scales)


def test_tri_graph_replay():
    tri, xg, ig, got = _run(CASES[0], torch.float32)
    out = torch.empty_like(got)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        tri.forward(xg, ig, torch.float32, out=out)
    s.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        tri.forward(xg, ig, torch.float32, out=out)
    out.zero_()
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, got)


def test_tri_graph_replay_one_accumulator():
    tri, xg, ig, got = _run(CASES[0], torch.bfloat16, fold=False)
    out = torch.empty_like(got)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        tri.forward(xg, ig, torch.bfloat16, out=out)
    s.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        tri.forward(xg, ig, torch.bfloat16, out=out)
    out.zero_()
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, got)


def test_tri_is_selected_for_tall_pair_groups():
    lib = _C.lib()
    pairs = 4 * 3 // 2
    cfg, w, table, ids, x = Hh.make_block(256, 4, 64, 128, 1000, 2, 16, seed=1)            # 0.5 / 0.5 x 2.0, 1.0
    blk = Hh.build_gpu_block(cfg, w, table, "bf16")
    assert lib.i64b_moe_tri_form(blk._h, pairs * _C.TRI_MIN_ROWS_ONEACC) == 2
    assert lib.i64b_moe_fuses_combine(blk._h, pairs * _C.TRI_MIN_ROWS_ONEACC) == 1
    assert lib.i64b_moe_tri_form(blk._h, pairs * _C.TRI_MIN_ROWS_ONEACC - 1) == 0
    cfg, w, table, ids, x = Hh.make_block(256, 4, 64, 128, 1000, 2, 16, seed=1, pw=0.95)   # not powers of two
    blk = Hh.build_gpu_block(cfg, w, table, "bf16")
    assert lib.i64b_moe_tri_form(blk._h, pairs * _C.TRI_MIN_ROWS_FOLD) == 1
    assert lib.i64b_moe_tri_form(blk._h, pairs * _C.TRI_MIN_ROWS_FOLD - 1) == 0


# ---- W8A8: int8 operands, int32 accumulators, per-segment dequant in the epilogues — bit-exact with the chain AND the oracle
W8_CASES = [
    dict(id="small", shape=(256, 4, 128, 128, 1000, 2), n=700),
    dict(id="zipf", shape=(256, 8, 128, 256, 5000, 2), n=3001, zipf=True),
    dict(id="collisions-95-05", shape=(512, 4, 256, 128, 1000, 2), n=901, collide=3, pw=0.95),
    dict(id="config5-8k", shape=(1024, 8, 512, 1280, 32000, 2), n=8192, sample=48),
]


def test_headline_step_against_oracle():
    """The step bench.py's headline is quoted on — BASELINE config 5 on one GPU: 65 536 tokens, H 1024, 8 experts x 512,
    shared 1280, top-2 — through the DEFAULT path (no debug flags: the form the library picks for this batch, which is the
    one that is timed), against the oracle on sampled rows at north_star's tolerance, plus the size-independent checks: every
    row finite and non-trivial, run-to-run identical, and equal to the GEMM2 + combine chain within the single-accumulator
    form's stated distance."""
    h, e, ie, s, v, k = 1024, 8, 512, 1280, 32000, 2
    n = 65536
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=5, pw=0.5, rs=2.0, ss=1.0)
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    blk = Hh.build_gpu_block(cfg, w, table, "bf16")
    assert blk.tri_form(n) == 2, "the headline batch should take the single-accumulator fused down projection"
    got = blk.forward(xg, ig, torch.float32)
    again = blk.forward(xg, ig, torch.float32)
    got16 = blk.forward(xg, ig, torch.bfloat16)          # bench.py's output dtype
    torch.cuda.synchronize()
    assert torch.equal(got, again)
    assert bool(torch.isfinite(got).all()) and int((got.abs().amax(dim=1) == 0).sum()) == 0
    assert bool(((got16.float() - got).abs() <= got.abs() * 2.0 ** -8 + 1e-30).all()), "bf16 output is not the fp32 output rounded once"
    rows = np.unique(np.concatenate([np.linspace(0, n - 1, 96).astype(int), np.arange(8), np.arange(n - 8, n)]))
    ref = O.moe_block_fp32(x[rows], ids[rows], table, w, cfg, inter_bf16=True)
    mx, mean = O.rel_errors(got.cpu().numpy()[rows], ref)
    assert mx <= Hh.TOL_MAX and mean <= Hh.TOL_MEAN, (mx, mean)
    chain = Hh.build_gpu_block(cfg, w, table, "bf16")
    chain.set_debug(_C.DEBUG_NO_FUSED | _C.DEBUG_NO_TRI)
    want = chain.forward(xg, ig, torch.float32)
    torch.cuda.synchronize()
    d, top = (got - want).abs().max().item(), want.abs().max().item()
    assert d <= top * 2e-5, (d, top)


@pytest.mark.parametrize("case", W8_CASES, ids=[c["id"] for c in W8_CASES])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_tri_w8a8_bit_exact(case, out_dtype):
    h, e, ie, s, v, k = case["shape"]
    n = case["n"]
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=5, zipf=case.get("zipf", False), pw=case.get("pw", 0.5))
    if case.get("collide"):
        table = np.array(table, copy=True)
        table[1, ::case["collide"]] = table[0, ::case["collide"]]
    wq = O.quantize_moe_weights(w)
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    chain = Hh.build_gpu_block(cfg, w, table, "int8", wq)
    chain.set_debug(_C.DEBUG_NO_TRI)
    tri = Hh.build_gpu_block(cfg, w, table, "int8", wq)
    tri.set_debug(_C.DEBUG_TRI)
    assert chain.tri_form(n) == 0 and tri.tri_form(n) == 1
    want = chain.forward(xg, ig, out_dtype)
    got = tri.forward(xg, ig, out_dtype)
    again = tri.forward(xg, ig, out_dtype)
    torch.cuda.synchronize()
    assert torch.equal(got, again), "not run-to-run identical"
    bad = (got != want).any(dim=1).nonzero().flatten()
    assert bad.numel() == 0, f"{bad.numel()} of {n} rows differ from the W8A8 chain, first {bad[:8].tolist()}"
    if out_dtype == torch.float32:
        sample = case.get("sample")
        rows = np.arange(n) if not sample else np.unique(np.linspace(0, n - 1, sample).astype(int))
        ref = O.moe_block_w8a8(x[rows], ids[rows], table, wq, cfg)
        assert np.array_equal(got.cpu().numpy()[rows], ref), "W8A8 fused down projection is not bit-exact with the oracle"


def test_tri_w8a8_preq_and_graph_replay():
    """x_preq rows (int8 + scale from the fused RMSNorm+quant) through the token-major path, captured and replayed."""
    h, e, ie, s, v, k = W8_CASES[0]["shape"]
    n = 1500
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=6)
    wq = O.quantize_moe_weights(w)
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    tri = Hh.build_gpu_block(cfg, w, table, "int8", wq)
    tri.set_debug(_C.DEBUG_TRI)
    want = tri.forward(xg, ig, torch.float32)
    xq, xs = O.quantize_rows_int8(x)
    got = tri.forward_preq(Hh.to_cuda(xq), Hh.to_cuda(xs), ig, torch.float32)
    assert torch.equal(got, want)
    out = torch.empty_like(want)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        tri.forward(xg, ig, torch.float32, out=out)
    st.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        tri.forward(xg, ig, torch.float32, out=out)
    out.zero_()
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    assert np.array_equal(want.cpu().numpy(), O.moe_block_w8a8(x, ids, table, wq, cfg))


@pytest.mark.skipif(os.environ.get("I64B_OVERLAP", "0") in ("", "0"), reason="opt-in path: run with I64B_OVERLAP=1 (measured slower, off by default)")
@pytest.mark.parametrize("form", ["chain", "one_acc"])
def test_dispatch_copy_overlap_is_bit_identical(form):
    """With I64B_OVERLAP=1, from 4096 tokens on, the dispatch row copies run on the layer's second stream under the shared
    expert's GEMM1 (I64B_DEBUG_NO_OVERLAP keeps them on the caller's stream): same kernels, identical output, also inside a
    captured graph."""
    n = 6000
    cfg, w, table, ids, x = Hh.make_block(256, 4, 64, 128, 1000, 2, n, seed=3)
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    flags = _C.DEBUG_NO_TRI if form == "chain" else _C.DEBUG_TRI
    serial = Hh.build_gpu_block(cfg, w, table, "bf16")
    serial.set_debug(flags | _C.DEBUG_NO_OVERLAP)
    over = Hh.build_gpu_block(cfg, w, table, "bf16")
    over.set_debug(flags)
    want = serial.forward(xg, ig, torch.float32)
    got = over.forward(xg, ig, torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    out = torch.empty_like(want)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        over.forward(xg, ig, torch.float32, out=out)
    st.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        over.forward(xg, ig, torch.float32, out=out)
        over.forward(xg, ig, torch.float32, out=out)
    out.zero_()
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    Hh.assert_close(got.cpu().numpy(), O.moe_block_fp32(x, ids, table, w, cfg, inter_bf16=True), "overlap")


@pytest.mark.parametrize("form", ["chain", "one_acc", "fold"])
@pytest.mark.parametrize("n", [20000, 33001])
def test_gemm1_row_gather_is_bit_identical(form, n):
    """Opt-in (I64B_DEBUG_GATHER; measured slower, off by default): blocks big enough for the CTA-pair GEMM1 make no
    dispatched copy, GEMM1's producer warp gathers the routed A rows from x (tile::gather4 through the dispatch indices).
    Same operands, same MMAs: identical output to the dispatched-copy path in every down-projection form, zipf ids included
    (ragged groups, empty tails)."""
    cfg, w, table, ids, x = Hh.make_block(256, 8, 128, 256, 5000, 2, n, seed=3, zipf=(n % 2 == 1))
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    flags = {"chain": _C.DEBUG_NO_TRI, "one_acc": _C.DEBUG_TRI, "fold": _C.DEBUG_TRI | _C.DEBUG_TRI_FOLD}[form]
    copy = Hh.build_gpu_block(cfg, w, table, "bf16")
    copy.set_debug(flags)
    gat = Hh.build_gpu_block(cfg, w, table, "bf16")
    gat.set_debug(flags | _C.DEBUG_GATHER)
    _C.launch_count_reset()
    want = copy.forward(xg, ig, torch.float32)
    torch.cuda.synchronize()
    l_copy = _C.launch_count()
    _C.launch_count_reset()
    got = gat.forward(xg, ig, torch.float32)
    again = gat.forward(xg, ig, torch.float32)
    torch.cuda.synchronize()
    assert _C.launch_count() == 2 * l_copy, "same number of launches (the dispatch launch writes indices only)"
    assert torch.equal(got, want) and torch.equal(got, again)
    rows = np.unique(np.linspace(0, n - 1, 96).astype(int))
    ref = O.moe_block_fp32(x[rows], ids[rows], table, w, cfg, inter_bf16=True)
    Hh.assert_close(got.cpu().numpy()[rows], ref, "gather")
