"""Generate golden vectors by running the REFERENCE itself (CPU, this container).

    PYTHONPATH=/root/reference PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

The reference (a pure-Python package) is imported read-only from
/root/reference; it does not travel to the GPU box, so the vectors are
committed next to this script (tests/golden/*.npz) and every consumer
(tests/, smoke) reads only those files.  Nothing from the reference's
source is copied: only inputs we generate and the outputs its functions
return are stored.
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("I64_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
sys.dont_write_bytecode = True
OUT = os.path.dirname(os.path.abspath(__file__))

from vllm_i64.kernels import i64_ops  # noqa: E402
from vllm_i64.kernels.kernel_loader import FallbackOps  # noqa: E402
from vllm_i64.kernels import fused_experts  # noqa: E402
from vllm_i64.core import quantization as Q  # noqa: E402
from vllm_i64.layers import integer_activations as IA  # noqa: E402
from vllm_i64.layers.moe import MoEExpert  # noqa: E402
from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP  # noqa: E402
from vllm_i64.layers.dense_mlp import DenseMLP  # noqa: E402
from vllm_i64.models.complexity_deep.model import DenseSwiGLUMLP  # noqa: E402

assert Q._INT_MM_CPU_OK, "reference would not take the torch._int_mm path on this CPU"
torch.set_num_threads(1)  # bit-stable fp32 reductions for the stored outputs


def npz(name, **kw):
    path = os.path.join(OUT, name)
    np.savez_compressed(path, **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v))
                                 for k, v in kw.items()})
    print(f"{name}: {os.path.getsize(path)/1024:.1f} KiB")


# ---- 1. routing KATs (reference tests/test_i64_pipeline.py:35-57, test_competitive_features.py:491-495)
kat = {}
kat["ids8"] = torch.arange(8, dtype=torch.int64)
kat["route8_e4"] = i64_ops.i64_route_tokens(kat["ids8"], 4)
kat["ids_b"] = torch.tensor([10, 11, 12, 13], dtype=torch.int64)
kat["route_b_e4"] = FallbackOps.route_tokens(kat["ids_b"], 4)
g = torch.Generator().manual_seed(42)
kat["ids_big"] = torch.randint(0, 100_000, (100_000,), generator=g, dtype=torch.int64)
for e in (2, 4, 8, 16):
    kat[f"route_big_e{e}"] = i64_ops.i64_route_tokens(kat["ids_big"], e)
npz("kat_routing.npz", **kat)

# ---- 2. the integer SiLU LUT (layers/integer_activations.py:28-34)
np.save(os.path.join(OUT, "silu_lut.npy"), IA._SILU_LUT.numpy().astype(np.int32))

# ---- 3. TokenRoutedMLP (the semantic spec): tiny config, fp32, CPU
H, E, IE, S, V, K, N = 128, 4, 32, 64, 512, 2, 53
torch.manual_seed(0)
mlp = TokenRoutedMLP(H, E * IE, E, V, shared_expert=True, shared_intermediate_size=S, top_k=K,
                     top_k_primary_weight=0.5, shared_output_scale=1.0, routed_output_scale=2.0).eval()
g = torch.Generator().manual_seed(7)
perm = torch.randperm(V, generator=g)
t0 = torch.empty(V, dtype=torch.long)
t0[perm] = torch.arange(V) % E
t1 = (t0 + 1 + (torch.arange(V) * 7 + 3) % (E - 1)) % E
table = torch.stack([t0, t1])
mlp.topk_token_to_expert.copy_(table)
mlp.token_to_expert.copy_(t0)
g = torch.Generator().manual_seed(1234)
ids = torch.randint(-3, V + 5, (N,), generator=g, dtype=torch.int64)   # includes out-of-range ids → clamp
g = torch.Generator().manual_seed(0)
x = torch.randn(N, H, generator=g)
with torch.no_grad():
    routes = mlp.route(ids, N, x.device)
    ef = [mlp.expert_forward(x, routes[r]) for r in range(K)]
    out = mlp(x, token_ids=ids)
    out_noids = mlp(x, token_ids=None)
    blk = dict(x=x, token_ids=ids, table=table, gate=mlp.gate_proj_w, up=mlp.up_proj_w, down=mlp.down_proj_w,
               shared_gate=mlp.shared_gate.weight, shared_up=mlp.shared_up.weight,
               shared_down=mlp.shared_down.weight, routes=routes, ef0=ef[0], ef1=ef[1], out=out,
               out_noids=out_noids, cfg=np.array([H, E, IE, S, V, K, N]),
               scales=np.array([0.5, 1.0, 2.0], dtype=np.float64))
    # cyclic default tables + top-2 property (reference tests/test_model.py:198-205)
    d = TokenRoutedMLP(H, E * IE, E, V, top_k=2)
    blk["default_token_to_expert"] = d.token_to_expert
    blk["default_topk"] = d.topk_token_to_expert
    # refactor-equivalence (tests/test_model.py:207-225)
    mlp.fuse_shared_gate_up()
    mlp.materialize_expert_linears()
    blk["out_fused"] = mlp(x, token_ids=ids)
    blk["w13_e0"] = mlp.expert_gate_up[0].weight
    blk["w2_e0"] = mlp.expert_down[0].weight
npz("tiny_block.npz", **blk)

# ---- 4. loader-table KAT (reference tests/test_loader.py:219-246): (arange%E, (arange*7+1)%E)
ar = torch.arange(V)
tbl = torch.stack([ar % E, (ar * 7 + 1) % E])
m2 = TokenRoutedMLP(H, E * IE, E, V, top_k=2)
m2.topk_token_to_expert.copy_(tbl)
qids = torch.tensor([0, 1, 5, 17, 100, V - 1, V + 10, -4])
npz("kat_loader_table.npz", table=tbl, ids=qids, routes=m2.route(qids, len(qids), torch.device("cpu")))

# ---- 5. W8A8 expert (layers/moe.py:114-122) with every intermediate
HW, IW, NW = 256, 128, 41
torch.manual_seed(3)
ex = MoEExpert(HW, IW).eval()
w13 = torch.cat([ex.w1.weight.data, ex.w3.weight.data], dim=0)
ex.w13_int8, ex.w13_scale = Q.quantize_int8(w13)
ex.w13_inter = IW
ex.w2_int8, ex.w2_scale = Q.quantize_int8(ex.w2.weight.data)
g = torch.Generator().manual_seed(11)
xw = torch.randn(NW, HW, generator=g) * 1.7
xw[5] = 0.0            # all-zero row → 1e-8 clamp
xw[6, 3] = 40.0        # large outlier → saturating LUT branch downstream
with torch.no_grad():
    x_q, x_s = Q.quantize_activations_int8(xw)
    gate, up = Q.int8_fused_gate_up_native(xw, ex.w13_int8, ex.w13_scale, IW)
    acc1 = torch._int_mm(x_q, ex.w13_int8.t().contiguous())
    inter = IA.silu_multiply_integer(gate, up)
    i_q, i_s = Q.quantize_activations_int8(inter)
    acc2 = torch._int_mm(i_q, ex.w2_int8.t().contiguous())
    y = ex(xw)
    y_lin = Q.int8_linear_native(inter, ex.w2_int8, ex.w2_scale)
    # LUT edge cases
    gq = torch.tensor([-2000, -1025, -1024, -1023, -1, 0, 1, 700, 1023, 1024, 1025, 5000], dtype=torch.int32)
    silu_edges = IA.silu_integer(gq)
    deq = Q.dequantize_int8(x_q, x_s)
npz("tiny_w8a8.npz", x=xw, w1=ex.w1.weight, w3=ex.w3.weight, w2=ex.w2.weight, w13_q=ex.w13_int8,
    w13_s=ex.w13_scale, w2_q=ex.w2_int8, w2_s=ex.w2_scale, x_q=x_q, x_s=x_s, acc1=acc1, gate=gate, up=up,
    inter=inter, inter_q=i_q, inter_s=i_s, acc2=acc2, y=y, y_lin=y_lin, silu_in=gq, silu_out=silu_edges,
    deq=deq)

# ---- 6. dense FFN (models/complexity_deep/model.py:243-276; layers/dense_mlp.py)
class _Cfg:
    hidden_size = 128
    intermediate_size = 192


torch.manual_seed(5)
dm = DenseSwiGLUMLP(_Cfg()).eval()
g = torch.Generator().manual_seed(21)
xd = torch.randn(19, 128, generator=g)
with torch.no_grad():
    yd = dm(xd)
    dn = dict(x=xd, gate=dm.gate_proj.weight, up=dm.up_proj.weight, down=dm.down_proj.weight, y=yd)
    dm.fuse_gate_up()
    dn["y_fused"] = dm(xd)
    dd = DenseMLP(128, 192).eval()
    dd.gate_proj.linear.weight.copy_(dn["gate"]); dd.up_proj.linear.weight.copy_(dn["up"])
    dd.down_proj.linear.weight.copy_(dn["down"])
    dn["y_densemlp"] = dd(xd)
    gu = torch.cat([dn["gate"], dn["up"]], dim=0)
    dd.gate_up_int8, dd.gate_up_scale = Q.quantize_int8(gu)
    dd.gate_up_inter = 192
    dd.down_int8, dd.down_scale = Q.quantize_int8(dn["down"])
    dn["y_int8"] = dd(xd)
npz("tiny_dense.npz", **dn)

# ---- 7. kernels package: scatter / gather / expert_forward / fused forward (kernels/i64_ops.py, fused_experts.py)
HK, EK, IK, NK = 64, 4, 32, 150
g = torch.Generator().manual_seed(31)
xk = torch.randn(NK, HK, generator=g)
idk = torch.randint(0, 10_000, (NK,), generator=g, dtype=torch.int64)
gu_w = torch.randn(EK, HK, 2 * IK, generator=g) * 0.1
dn_w = torch.randn(EK, IK, HK, generator=g) * 0.1
with torch.no_grad():
    eid = i64_ops.i64_route_tokens(idk, EK)
    sc, sidx, offs, cnts = i64_ops.i64_scatter(xk, eid, EK)
    stable_idx = eid.long().argsort(stable=True)
    ef_out = i64_ops.i64_expert_forward(sc, gu_w, dn_w, offs, cnts, EK)
    gathered = i64_ops.i64_gather(ef_out, sidx.long())
    full_chunked = i64_ops.i64_full_pipeline(idk, xk, gu_w, dn_w, EK)                 # N > 64 → chunked
    full_bmm = fused_experts.fused_token_routed_forward(xk[:40], gu_w, dn_w, eid[:40].long(), EK, IK)
    qe = Q.quantize_experts(gu_w, dn_w, Q.QuantConfig(method="int8"))
    full_int8 = fused_experts.fused_token_routed_forward_int8(
        xk, qe["gate_up_int8"], qe["gate_up_scale"], qe["down_int8"], qe["down_scale"], eid.long(), EK, IK)
    fr = FallbackOps.fused_route_scatter(idk, xk, EK)
    sh = FallbackOps.silu_hadamard(xk @ gu_w[0], IK)
npz("tiny_kernels.npz", x=xk, token_ids=idk, gate_up=gu_w, down=dn_w, expert_ids=eid, scattered=sc,
    scatter_indices=sidx, stable_argsort=stable_idx, offsets=offs, counts=cnts, expert_out=ef_out,
    gathered=gathered, full_chunked=full_chunked, full_bmm=full_bmm, gu_q=qe["gate_up_int8"],
    gu_s=qe["gate_up_scale"], d_q=qe["down_int8"], d_s=qe["down_scale"], full_int8=full_int8,
    frs_expert_ids=fr[0], frs_scattered=fr[1], frs_map=fr[2], frs_offsets=fr[3], frs_counts=fr[4],
    silu_hadamard=sh)
print("done")
