"""Token-major fused down projection (i64b_launch_gemm_tri): GEMM2 of both routes and of the shared expert + the weighted
combine in ONE kernel, three TMEM accumulators per tile, no fp32 per-assignment rows in memory.  It applies the same fp32
operations in the same order as GEMM2 + combine_kernel, so the bar is BIT IDENTITY with that chain (and the chain's own
bar against the oracle).  Cases: small and config-5 geometry, no shared expert, skewed ids (empty / tiny / huge pair
groups), token counts that are not a multiple of any tile, both output dtypes, unequal route weights, graph replay."""
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
    dict(id="weights-95-05", shape=(256, 4, 64, 128, 1000, 2), n=1300, pw=0.95),
    dict(id="config5-8k", shape=(1024, 8, 512, 1280, 32000, 2), n=8192, sample=64),
    dict(id="config5-33k", shape=(1024, 8, 512, 1280, 32000, 2), n=33000, sample=32),
]


def _run(case, out_dtype):
    h, e, ie, s, v, k = case["shape"]
    n = case["n"]
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=11, zipf=case.get("zipf", False),
                                          pw=case.get("pw", 0.5))
    xg, ig = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    chain = Hh.build_gpu_block(cfg, w, table, "bf16")
    chain.set_debug(_C.DEBUG_NO_FUSED | _C.DEBUG_NO_TRI)
    tri = Hh.build_gpu_block(cfg, w, table, "bf16")
    tri.set_debug(_C.DEBUG_NO_FUSED | _C.DEBUG_TRI)
    want = chain.forward(xg, ig, out_dtype)
    got = tri.forward(xg, ig, out_dtype)
    again = tri.forward(xg, ig, out_dtype)
    torch.cuda.synchronize()
    assert torch.equal(got, again), "not run-to-run identical"
    bad = (got != want).any(dim=1).nonzero().flatten()
    assert bad.numel() == 0, f"{bad.numel()} of {n} rows differ from the GEMM2 + combine chain, first {bad[:8].tolist()}"
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
    _run(case, out_dtype)


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


def test_tri_is_selected_for_tall_pair_groups():
    cfg, w, table, ids, x = Hh.make_block(256, 4, 64, 128, 1000, 2, 16, seed=1)
    blk = Hh.build_gpu_block(cfg, w, table, "bf16")
    lib = _C.lib()
    assert lib.i64b_moe_fuses_combine(blk._h, 4 * 3 * 768) == 1          # auto: rows per pair >= 768
    assert lib.i64b_moe_fuses_combine(blk._h, 4 * 3 * 768 - 1) == 0
