"""``torch.ops.vllm_i64.*`` (kernels/torch_ops.py): the registration the reference makes with TORCH_LIBRARY(vllm_i64)
(csrc/i64_ops_binding.cu:400-407) plus the whole-layer op that lets ``torch.compile`` trace through a block
(layers/token_routed_i64/fused_i64_moe.py:123-142).  CPU tests: names / schemas / fake impls / tracing; GPU tests: every
op against the oracle through ``torch.ops``, ``torch.library.opcheck``, and a compiled module without graph breaks."""
import copy
import os
import re

import numpy as np
import pytest
import torch

import i64_test_helpers as Hh
from oracle import i64_oracle as O

from vllm_i64.kernels import torch_ops
from vllm_i64.layers.dense_mlp import DenseMLP
from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP

ops = torch.ops.vllm_i64
gpu = pytest.mark.gpu
needs_cuda = pytest.mark.skipif(not torch.cuda.is_available(), reason="CUDA device required")


def test_six_reference_ops_are_registered_with_reference_signatures():
    for name in torch_ops.REFERENCE_OPS:
        assert hasattr(ops, name), name
        schema = str(getattr(ops, name).default._schema)
        assert schema == f"vllm_i64::{name}{torch_ops.SCHEMAS[name]}"
    # argument names / order of the reference's C++ ops, when the reference tree is at hand (build container only)
    src = "/root/reference/csrc/i64_ops_binding.cu"
    if not os.path.exists(src):
        return
    text = open(src).read()
    for name in torch_ops.REFERENCE_OPS:
        m = re.search(r"\b%s\s*\(([^)]*)\)\s*\{" % name, text)
        assert m, name
        ref_args = [a.strip().split()[-1].lstrip("&*") for a in m.group(1).split(",") if a.strip()]
        ours = [a.name for a in getattr(ops, name).default._schema.arguments]
        assert ours == ref_args, (name, ours, ref_args)


def test_fake_impls_give_shapes_and_dtypes():
    x = torch.empty(6, 64, device="meta", dtype=torch.bfloat16)
    ids = torch.empty(6, device="meta", dtype=torch.int64)
    assert ops.route_tokens(ids, 8).dtype == torch.int32 and ops.route_tokens(ids, 8).shape == (6,)
    assert ops.silu_hadamard(torch.empty(6, 128, device="meta"), 64).shape == (6, 64)
    sc, idx, off, cnt = ops.scatter_by_expert(x, ids.to(torch.int32), 8)
    assert sc.shape == x.shape and sc.dtype == x.dtype and idx.shape == (6,) and off.shape == (8,) and cnt.shape == (8,)
    assert all(t.dtype == torch.int32 for t in (idx, off, cnt))
    assert ops.gather_by_expert(x, idx).shape == x.shape
    assert len(ops.atomic_scatter(x, ids.to(torch.int32), 8)) == 4
    e, sc, mp, off, cnt = ops.fused_route_scatter(ids, x, 8)
    assert e.dtype == torch.int32 and sc.shape == x.shape and off.shape == (8,)
    h, y, q, s = ops.add_rmsnorm(x, x, torch.empty(64, device="meta", dtype=torch.bfloat16), 1e-6, True)
    assert q.dtype == torch.int8 and q.shape == x.shape and s.shape == (6, 1)
    assert ops.token_routed_mlp(x, ids, None, None, 1).shape == x.shape


def test_cpu_tensors_raise():
    with pytest.raises(NotImplementedError):
        ops.route_tokens(torch.zeros(4, dtype=torch.int64), 8)


def test_layer_ids_are_unique_and_survive_copies():
    a = TokenRoutedMLP(64, 8 * 32, 8, 100, top_k=2)
    b = copy.deepcopy(a)
    d = DenseMLP(64, 128)
    assert len({a._i64_layer_id, b._i64_layer_id, d._i64_layer_id}) == 3
    assert torch_ops._LAYERS[b._i64_layer_id] is b and torch_ops._LAYERS[a._i64_layer_id] is a


def test_compile_traces_a_block_as_one_node():
    """fullgraph=True: dynamo must not break at the block; the graph holds exactly one vllm_i64::token_routed_mlp node."""
    m = TokenRoutedMLP(64, 8 * 32, 8, 100, shared_expert=True, shared_intermediate_size=64, top_k=2)
    graphs = []

    def backend(gm, example_inputs):
        graphs.append(gm)
        return lambda *args: (torch.zeros_like(args[0]),)

    fn = torch.compile(lambda x, ids: m(x, token_ids=ids) + 1.0, backend=backend, fullgraph=True)
    x = torch.randn(5, 64)
    fn(x, torch.arange(5))
    assert len(graphs) == 1
    targets = [n.target for n in graphs[0].graph.nodes if n.op == "call_function"]
    assert sum(t in (ops.token_routed_mlp, ops.token_routed_mlp.default) for t in targets) == 1, targets


# ------------------------------------------------------------------------------------------------------------- GPU
@gpu
@needs_cuda
def test_ops_match_the_oracle_through_torch_ops():
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(3)
    n, h, e, inter = 333, 256, 8, 96
    ids = rng.integers(0, 50000, n).astype(np.int64)
    hidden = rng.standard_normal((n, h)).astype(np.float32)
    t_ids, t_hidden = torch.from_numpy(ids).to(dev), torch.from_numpy(hidden).to(dev)
    eids = ops.route_tokens(t_ids, e)
    assert np.array_equal(eids.cpu().numpy(), (ids & (e - 1)).astype(np.int32))
    sc, idx, off, cnt = ops.scatter_by_expert(t_hidden, eids, e)
    order = np.argsort(eids.cpu().numpy(), kind="stable")
    pos = np.empty(n, np.int32)                     # scatter_indices[i] = row of token i in the expert-sorted buffer
    pos[order] = np.arange(n, dtype=np.int32)
    assert np.array_equal(idx.cpu().numpy(), pos)
    assert np.array_equal(sc.cpu().numpy(), hidden[order])
    assert np.array_equal(cnt.cpu().numpy(), np.bincount(eids.cpu().numpy(), minlength=e))
    assert np.array_equal(off.cpu().numpy(), np.concatenate([[0], np.cumsum(cnt.cpu().numpy())[:-1]]))
    assert torch.equal(ops.gather_by_expert(sc, idx), t_hidden)          # round trip
    e2, sc2, mp2, off2, cnt2 = ops.fused_route_scatter(t_ids, t_hidden, e)
    assert torch.equal(e2, eids) and torch.equal(sc2, sc) and torch.equal(mp2, idx) and torch.equal(cnt2, cnt)
    gu = rng.standard_normal((n, 2 * inter)).astype(np.float32)
    got = ops.silu_hadamard(torch.from_numpy(gu).to(dev), inter).cpu().numpy()
    want = O.silu_hadamard(gu, inter) if hasattr(O, "silu_hadamard") else (gu[:, :inter] / (1 + np.exp(-gu[:, :inter]))) * gu[:, inter:]
    assert np.abs(got - want).max() <= 1e-5 * max(1.0, np.abs(want).max())


@gpu
@needs_cuda
def test_opcheck():
    dev = torch.device("cuda", 0)
    ids = torch.randint(0, 1000, (64,), device=dev)
    x = torch.randn(64, 128, device=dev)
    eids = ops.route_tokens(ids, 8)
    kinds = ("test_schema", "test_faketensor")
    torch.library.opcheck(ops.route_tokens.default, (ids, 8), test_utils=kinds)
    torch.library.opcheck(ops.silu_hadamard.default, (x, 64), test_utils=kinds)
    torch.library.opcheck(ops.scatter_by_expert.default, (x, eids, 8), test_utils=kinds)
    torch.library.opcheck(ops.gather_by_expert.default, (x, eids % 64), test_utils=kinds)
    torch.library.opcheck(ops.fused_route_scatter.default, (ids, x, 8), test_utils=kinds)
    xb = x.to(torch.bfloat16)
    w = torch.ones(128, device=dev, dtype=torch.bfloat16)
    torch.library.opcheck(ops.add_rmsnorm.default, (xb, xb, w, 1e-6, True), test_utils=kinds)


@gpu
@needs_cuda
@pytest.mark.parametrize("quant", ["none", "int8"])
def test_compiled_block_equals_eager(quant):
    """torch.compile (aot_eager backend: tracing + functionalisation, no code generation) over norm -> block -> residual:
    no graph break, same bits as the eager call."""
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    m = TokenRoutedMLP(256, 8 * 64, 8, 1000, shared_expert=True, shared_intermediate_size=128, top_k=2).to(dev, torch.bfloat16)
    if quant == "int8":
        m.quantize_w8a8()
    x = torch.randn(77, 256, device=dev, dtype=torch.bfloat16)
    ids = torch.randint(0, 1000, (77,), device=dev)

    def f(x, ids):
        return x + m(x * 0.5, token_ids=ids)
    want = f(x, ids)
    got = torch.compile(f, backend="aot_eager", fullgraph=True)(x, ids)
    torch.cuda.synchronize()
    assert torch.equal(got, want)
