"""HostStepPipeline (kernels/block.py): an engine step over pinned HOST buffers, cut into micro-batches whose PCIe copies run
under the kernels of their neighbours.  Tokens of an FFN block are independent, so every cut must give the bits of the
one-shot device call; also under CUDA-graph replay and with chained layers."""
import numpy as np
import pytest
import torch

import i64_test_helpers as Hh

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("CUDA device required", allow_module_level=True)

from vllm_i64 import _C
from vllm_i64.kernels.block import HostStepPipeline


def _layers(mode, n, count=3):
    out = []
    for layer in range(count):
        cfg, w, table, ids, x = Hh.make_block(256, 4, 64, 128, 1000, 2, n, seed=20 + layer, layer=layer)
        wq = None
        if mode == "int8":
            from oracle import i64_oracle as O
            wq = O.quantize_moe_weights(w)
        blk = Hh.build_gpu_block(cfg, w, table, mode, wq)
        blk.set_debug(_C.DEBUG_NO_FUSED)       # one arithmetic path whatever the micro-batch size (the decode kernel splits K differently)
        out.append(blk)
    return out, ids, x


@pytest.mark.parametrize("mode", ["bf16", "int8"])
@pytest.mark.parametrize("chunks", [1, 2, 3, 5, (1, 2, 3, 2, 1)], ids=["1", "2", "3", "5", "uneven"])
@pytest.mark.parametrize("chain", [False, True])
def test_pipeline_equals_one_shot(mode, chunks, chain):
    n = 1537
    layers, ids, x = _layers(mode, n)
    x_host = torch.from_numpy(x).to(torch.bfloat16).pin_memory()
    ids_host = torch.from_numpy(ids).pin_memory()
    out_host = torch.empty(n, 256, dtype=torch.bfloat16).pin_memory()
    xd, idd = x_host.cuda(), ids_host.cuda()
    want = None
    src = xd
    for blk in layers:
        want = blk.forward(src, idd, torch.bfloat16)
        if chain:
            src = want
    pipe = HostStepPipeline(layers, n, chunks=chunks)
    pipe.run(x_host, ids_host, out_host, chain=chain)
    torch.cuda.synchronize()
    assert torch.equal(out_host, want.cpu())
    # fewer tokens than the capacity, replayed from a CUDA graph
    m = n - 300
    want_m = None
    src = xd[:m]
    for blk in layers:
        want_m = blk.forward(src, idd[:m], torch.bfloat16)
        if chain:
            src = want_m
    s = torch.cuda.Stream()
    xh, ih, oh = x_host[:m], ids_host[:m], out_host[:m]
    with torch.cuda.stream(s):
        pipe.run(xh, ih, oh, chain=chain)
    s.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        pipe.run(xh, ih, oh, chain=chain)
    out_host.zero_()
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out_host[:m], want_m.cpu())


@pytest.mark.parametrize("chunks", [3, (1, 2, 3, 2, 1)], ids=["3", "uneven"])
def test_streaming_steps_equal_one_shot(chunks):
    """run_stream: consecutive steps with DIFFERENT inputs and output buffers, nothing joined between them (the next step's
    copy-in overlaps this step's kernels, slot events carry over), one drain at the end: every step's output is the one-shot
    block's, bit for bit; a second round over the same buffers too."""
    n = 1537
    layers, ids, x = _layers("bf16", n)
    rng = np.random.RandomState(3)
    steps = []
    for k in range(4):
        xk = torch.from_numpy(np.roll(x, k * 37, axis=0)).to(torch.bfloat16).pin_memory()
        ik = torch.from_numpy(np.roll(ids, k * 37)).pin_memory()
        m = n - 200 * k                                   # steps of different lengths
        steps.append((xk[:m], ik[:m], torch.empty(m, 256, dtype=torch.bfloat16).pin_memory()))
    want = []
    for xk, ik, _ in steps:
        w = None
        for blk in layers:
            w = blk.forward(xk.cuda(), ik.cuda(), torch.bfloat16)
        want.append(w.cpu())
    pipe = HostStepPipeline(layers, n, chunks=chunks)
    for rnd in range(2):
        for _, _, ok in steps:
            ok.zero_()
        for xk, ik, ok in steps:
            pipe.run_stream(xk, ik, ok)
        pipe.drain()
        torch.cuda.synchronize()
        for (_, _, ok), w in zip(steps, want):
            assert torch.equal(ok, w), f"round {rnd}"
    # and the self-contained form still works after streaming
    pipe.run(steps[0][0], steps[0][1], steps[0][2])
    torch.cuda.synchronize()
    assert torch.equal(steps[0][2], want[0])
    # ... also when the streamed steps were NOT drained first: run() orders itself behind their copy-outs, which still read the
    # slots' output buffers, instead of overwriting them
    for _ in range(3):
        for _, _, ok in steps:
            ok.zero_()
        for xk, ik, ok in steps[:3]:
            pipe.run_stream(xk, ik, ok)
        pipe.run(steps[3][0], steps[3][1], steps[3][2])
        torch.cuda.synchronize()
        for (_, _, ok), w in zip(steps, want):
            assert torch.equal(ok, w), "run() after undrained run_stream()"
