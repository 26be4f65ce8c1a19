"""Worker of tests/test_dropin_reference.py::test_engine_wiring_on_gpu (fresh interpreter, this repo's package first on the
path, the reference copy second): the reference's OWN post-load quantisation passes, CUDA-graph runner and model class driven
over the sm_100a modules.  Reads the checkpoints the "ref" run of dropin_worker.py wrote."""
import json
import os
import sys

import torch

workdir = sys.argv[1]
import vllm_i64  # noqa: E402

hooked = vllm_i64.patch_reference_models()
from vllm_i64 import _C  # noqa: E402
from vllm_i64.core import loader  # noqa: E402  (the reference's file)
from vllm_i64.core.cuda_graph import CUDAGraphRunner  # noqa: E402  (the reference's file)
from vllm_i64.layers.dense_mlp import DenseMLP  # noqa: E402
from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP  # noqa: E402
from vllm_i64.models.complexity_deep.model import ComplexityDeepModel  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.argv = [sys.argv[0], "config-only", workdir, "none", "fp32"]
info = {"hooked": hooked}
dev = "cuda:0"
g = torch.Generator().manual_seed(0)
ids = torch.randint(0, 256, (2, 24), generator=g).to(dev)


def config(moe=True):       # the same tiny configs as dropin_worker.py
    from vllm_i64.models.complexity_deep.config import ComplexityDeepConfig
    return ComplexityDeepConfig(
        hidden_size=128, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
        intermediate_size=512, num_experts=4 if moe else 1, vocab_size=256, max_position_embeddings=128,
        use_token_routed_mlp=moe, shared_expert=moe, shared_intermediate_size=64 if moe else 0, top_k=2 if moe else 1,
        top_k_primary_weight=0.5, shared_output_scale=1.0, routed_output_scale=2.0, use_mu_guidance=True, use_qk_norm=True)


def build(kind):
    torch.manual_seed(1 if kind == "moe" else 2)
    m = ComplexityDeepModel(config(kind == "moe")).eval()
    loader.load_checkpoint(m, os.path.join(workdir, f"{kind}.pt"), dtype=torch.float32, device="cpu")
    return m


def rel(a, b):
    d = (a.float() - b.float()).abs()
    return float(d.max() / b.float().abs().max()), float(d.mean() / b.float().abs().mean())


with torch.no_grad():
    # ---- f3: --quantization int8 reaches the W8A8 kernels through the reference's own passes (core/loader.py:627-642)
    moe = build("moe")
    base = moe.to(dev).to(torch.bfloat16)(ids)
    moe = build("moe")                                   # the loader quantises BEFORE it moves the model: host weights
    loader._quantize_experts(moe, "int8")
    loader._quantize_dense_mlp(moe, "int8")
    mlps = [m for m in moe.modules() if isinstance(m, TokenRoutedMLP)]
    info["moe_quantization"] = sorted({m.quantization for m in mlps})
    _C.launch_count_reset()
    q = moe.to(dev).to(torch.bfloat16)(ids)
    info["moe_int8_launches"] = _C.launch_count()
    info["moe_int8_vs_bf16"] = rel(q, base)
    info["moe_block_modes"] = sorted({int(m._block.mode) for m in mlps if m._block is not None})
    # ---- f3: device-side expert histogram for /v1/experts (api/_admin.py:464-562)
    load = vllm_i64.collect_expert_load(moe)
    info["expert_load_layers"] = len(load)
    info["expert_load_row_sums"] = sorted({sum(row) for counts in load.values() for row in counts})
    # ---- DenseMLP (layers/dense_mlp.py; no registered model instantiates it, the passes are written for it): the buffers
    # _quantize_dense_mlp / _quantize_dense_mlp_fp8 register are consumed as they are
    class Holder(torch.nn.Module):
        def __init__(self):
            super().__init__()
            torch.manual_seed(7)
            self.mlp = DenseMLP(128, 512)
    xd = torch.randn(40, 128, generator=torch.Generator().manual_seed(8)).to(dev)
    dbase = Holder().to(dev)(xd) if False else Holder().to(dev).mlp(xd)
    for method in ("int8", "fp8"):
        dense = Holder()                                  # host weights, as in the loader
        if method == "int8":
            loader._quantize_dense_mlp(dense, "int8")
        else:
            loader._quantize_dense_mlp_fp8(dense)
        dm = [m for m in dense.modules() if isinstance(m, DenseMLP)]
        out = dense.to(dev).mlp(xd)
        info[f"dense_{method}_modules"] = len(dm)
        info[f"dense_{method}_quantization"] = sorted({m.quantization for m in dm})
        info[f"dense_{method}_vs_bf16"] = rel(out, dbase)
    # ---- f2: the reference's CUDAGraphRunner (core/cuda_graph.py:21-153) captures the block for its common batch sizes
    moe = build("moe").to(dev).to(torch.bfloat16)
    mlp = moe.layers[0].mlp
    xbuf = torch.randn(64, 128, device=dev, generator=torch.Generator(device=dev).manual_seed(5)).to(torch.bfloat16)

    def forward_fn(token_ids, positions, expert_ids):
        return mlp(xbuf[: token_ids.shape[0]], token_ids)

    runner = CUDAGraphRunner(forward_fn, max_batch_size=64, device=dev)
    runner.capture_common_sizes()
    info["graph_sizes"] = sorted(runner._captured_sizes)
    ok = []
    for bs in (1, 5, 33, 64):
        tok = torch.randint(0, 256, (bs,), device=dev, generator=torch.Generator(device=dev).manual_seed(bs))
        pos = torch.zeros(bs, dtype=torch.int32, device=dev)
        eid = torch.zeros(bs, dtype=torch.int32, device=dev)
        got = runner.run(tok, pos, eid)
        gs = runner._find_best_size(bs)
        padded = torch.zeros(gs, dtype=torch.int64, device=dev)
        padded[:bs] = tok
        want = mlp(xbuf[:gs], padded)[:bs]
        ok.append(bool(torch.equal(got, want)))
    info["graph_replay_equals_eager"] = ok
print("WIRING_JSON " + json.dumps(info))
