"""Fused decode kernel (csrc/i64b_fused.cu): shapes at the edges of what it accepts, and its opt-in weight prefetch.

The block-level parity of the fused kernel against the oracle lives in test_gpu_parity.py (BLOCK_CASES with N <= 1024);
here: the dense FFN of BASELINE config 4 at decode size (8 split-K slabs of the down-projection, 64 GEMM1 tiles per
m-tile: both at the kernel's limits), a shape just past the limit (must take the unfused chain and still agree), and the
cross-layer L2 prefetch switch (a separate process, because the library reads the environment once)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import i64_test_helpers as Hh
from oracle import i64_oracle as O

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("CUDA device required", allow_module_level=True)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _dense(h, s, n, seed):
    rng = np.random.RandomState(seed)
    b = 1.0 / np.sqrt(h)
    gate = Hh.bf16_round(rng.uniform(-b, b, (s, h)).astype(np.float32))
    up = Hh.bf16_round(rng.uniform(-b, b, (s, h)).astype(np.float32))
    down = Hh.bf16_round(rng.uniform(-1.0 / np.sqrt(s), 1.0 / np.sqrt(s), (h, s)).astype(np.float32))
    x = Hh.bf16_round(rng.randn(n, h).astype(np.float32))
    return gate, up, down, x


@pytest.mark.parametrize("h,s,n,fused", [(1024, 4096, 256, True),     # dense-306 decode (config 4): 8 slabs, 64 tiles
                                         (256, 4096 + 64, 100, False),  # one tile past the fused kernel's limit
                                         (256, 1536, 1024, True)])     # largest fused batch, 8 m-tiles
def test_dense_decode_sizes(h, s, n, fused):
    from vllm_i64.layers.dense_mlp import DenseSwiGLUMLP
    gate, up, down, x = _dense(h, s, n, seed=s + n)

    class Cfg:
        hidden_size, intermediate_size = h, s

    m = DenseSwiGLUMLP(Cfg())
    m.load_state_dict({"gate_proj.weight": torch.from_numpy(gate), "up_proj.weight": torch.from_numpy(up),
                       "down_proj.weight": torch.from_numpy(down)})
    m = m.cuda()
    y = m(Hh.to_cuda(x))
    blk = getattr(m, "_block", None)
    if blk is not None:
        assert bool(blk.uses_fused(n)) == fused
    Hh.assert_close(y.float().cpu().numpy(), O.dense_swiglu_fp32(x, gate, up, down, True), f"dense decode {h}x{s} n={n}")
    assert torch.equal(m(Hh.to_cuda(x)), y)                    # run-to-run bit-identical


_PREFETCH_SCRIPT = r"""
import sys, hashlib
sys.path[:0] = [{root!r}, {root!r} + "/vllm-i64_b200", {root!r} + "/tests"]
import numpy as np, torch
import i64_test_helpers as Hh
outs = []
blocks = []
for layer in range(3):
    cfg, w, table, ids, x = Hh.make_block(256, 8, 64, 128, 1000, 2, 200, seed=layer, layer=layer)
    blocks.append((Hh.build_gpu_block(cfg, w, table, "bf16"), ids, x))
for i, (b, _, _) in enumerate(blocks):
    b.set_next(blocks[(i + 1) % 3][0])
h = hashlib.sha256()
for rep in range(2):
    for b, ids, x in blocks:
        assert b.uses_fused(200)
        out = b.forward(Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids), torch.float32)
        h.update(out.cpu().numpy().tobytes())
print("DIGEST", h.hexdigest())
"""


def _run(env_value):
    env = dict(os.environ, PYTHONDONTWRITEBYTECODE="1")
    env.pop("I64B_PREFETCH", None)
    if env_value is not None:
        env["I64B_PREFETCH"] = env_value
    r = subprocess.run([sys.executable, "-c", _PREFETCH_SCRIPT.format(root=ROOT)], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return [ln for ln in r.stdout.splitlines() if ln.startswith("DIGEST")][-1]


def test_prefetch_switch_does_not_change_results():
    """i64b_moe_set_next only moves bytes into L2: outputs with the prefetch at kernel entry (1), after the layer's own
    weight stream (2) and off (default) are bit-identical."""
    base = _run(None)
    assert _run("1") == base
    assert _run("2") == base


@pytest.mark.parametrize("case", [(1024, 8, 512, 1280, 32000, 2, 256),     # config 2
                                  (256, 4, 64, 96, 700, 3, 700),          # several 128-row passes per expert
                                  (512, 16, 48, 0, 5000, 1, 1000)])       # no shared expert, ragged widths
def test_back_to_back_launches_are_bit_identical(case):
    """Soak: the kernel hands work between CTAs through device counters that the last CTA resets, stages rows through
    shared memory and overlaps consecutive launches (PDL).  200 graph replays of four chained layers must reproduce the
    first replay bit for bit — a race in any of that shows up as a differing output."""
    h, e, ie, s, v, k, n = case
    blocks, xs = [], None
    for layer in range(4):
        cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=layer + 3, layer=layer)
        blocks.append(Hh.build_gpu_block(cfg, w, table, "bf16"))
        if xs is None:
            xs, idt = Hh.to_cuda(x, torch.bfloat16), Hh.to_cuda(ids)
    assert blocks[0].uses_fused(n)
    outs = [torch.empty(n, h, dtype=torch.bfloat16, device="cuda") for _ in range(4)]
    stream = torch.cuda.Stream()

    def run():
        for i, b in enumerate(blocks):                     # layer i reads the previous layer's output
            b.forward(xs if i == 0 else outs[i - 1], idt, torch.bfloat16, out=outs[i])

    with torch.cuda.stream(stream):
        run()
    stream.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        run()
    with torch.cuda.stream(stream):
        g.replay()
        stream.synchronize()
        first = [o.clone() for o in outs]
        for rep in range(200):
            g.replay()
            if rep % 20 == 19:
                stream.synchronize()
                for a, b in zip(first, outs):
                    assert torch.equal(a, b), f"replay {rep} differs"
    stream.synchronize()
    assert all(torch.isfinite(o.float()).all() for o in outs)
