"""Worker of tests/test_dropin_reference.py (run in a fresh interpreter with a controlled PYTHONPATH).

  mode "ref"     : only the reference package (oracle/_ref) is importable.  Builds the reference's ComplexityDeepModel
                   (tiny tr-moe config: top-2 routing, shared expert, layer-specific tables), saves its state dict as a
                   checkpoint and its logits on a fixed batch.
  mode "overlay" : this repo's vllm_i64 comes FIRST on the path, the reference second — the package overlays itself on
                   the reference (vllm_i64/_overlay.py).  Builds the SAME reference model class (whose MLPs now resolve to the
                   sm_100a modules), loads the checkpoint through the reference's own core/loader.load_checkpoint and writes
                   its logits.  On a GPU the logits are computed on cuda (the product path); on a CPU-only box the worker
                   stops after the structural checks.
"""
import json
import os
import sys

import torch

mode, workdir, device, dtype_name = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4]
dtype = {"fp32": torch.float32, "bf16": torch.bfloat16}[dtype_name]

from vllm_i64.models.complexity_deep.config import ComplexityDeepConfig  # noqa: E402
from vllm_i64.models.complexity_deep.model import ComplexityDeepModel  # noqa: E402
import vllm_i64.layers.token_routed_mlp as trm  # noqa: E402

if mode == "overlay":
    import vllm_i64
    vllm_i64.patch_reference_models()          # DenseSwiGLUMLP is defined inside the reference's model file


def config(moe=True):
    return ComplexityDeepConfig(
        hidden_size=128, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
        intermediate_size=512, num_experts=4 if moe else 1, vocab_size=256, max_position_embeddings=128,
        use_token_routed_mlp=moe, shared_expert=moe, shared_intermediate_size=64 if moe else 0, top_k=2 if moe else 1,
        top_k_primary_weight=0.5, shared_output_scale=1.0, routed_output_scale=2.0, use_mu_guidance=True, use_qk_norm=True)


info = {"mode": mode, "trm_file": os.path.realpath(trm.__file__)}
g = torch.Generator().manual_seed(0)
ids = torch.randint(0, 256, (2, 24), generator=g)
for kind in ("moe", "dense"):
    torch.manual_seed(1 if kind == "moe" else 2)
    model = ComplexityDeepModel(config(kind == "moe")).eval()
    ckpt = os.path.join(workdir, f"{kind}.pt")
    if mode == "ref":
        if kind == "moe":
            for li, layer in enumerate(model.layers):        # layer-specific top-2 tables (the checkpoint ships them)
                v = torch.arange(256)
                t0 = (v * 7 + li * 3) % 4
                layer.mlp.topk_token_to_expert.copy_(torch.stack([t0, (t0 + 1 + (v + li) % 3) % 4]))
                layer.mlp.token_to_expert.copy_(t0)
        torch.save(model.state_dict(), ckpt)
    else:
        from vllm_i64.core.loader import load_checkpoint
        stats = load_checkpoint(model, ckpt, dtype=torch.float32, device="cpu")
        info[f"{kind}_load_stats"] = {k: (v if isinstance(v, (int, float, str)) else len(v)) for k, v in stats.items()}
    info[f"{kind}_mlp_class"] = f"{type(model.layers[0].mlp).__module__}.{type(model.layers[0].mlp).__qualname__}"
    info[f"{kind}_mlp_file"] = os.path.realpath(sys.modules[type(model.layers[0].mlp).__module__].__file__)
    info[f"{kind}_keys"] = sorted(model.state_dict().keys())
    if device != "none":
        model = model.to(device).to(dtype)
        with torch.no_grad():
            logits = model(ids.to(device))
        torch.save(logits.float().cpu(), os.path.join(workdir, f"{kind}_{mode}_logits.pt"))
if mode == "overlay" and device.startswith("cuda"):
    from vllm_i64 import _C
    info["launches"] = _C.launch_count()
print("DROPIN_JSON " + json.dumps(info))
