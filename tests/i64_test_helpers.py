"""Shared builders for the parity tests: seeded synthetic inputs in numpy, usable by the oracle
(CPU) and the CUDA path (after upload) alike."""
from __future__ import annotations

import numpy as np

from oracle import i64_oracle as O


def bf16_round(a: np.ndarray) -> np.ndarray:
    """Round fp32 to the nearest bf16 (ties to even), returned as fp32 — the values the GPU sees."""
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(torch.bfloat16).float().numpy()


def make_block(h, e, ie, s, v, k, n, seed=0, layer=0, zipf=False, pw=0.5, rs=2.0, ss=1.0, scale=None):
    """Random-init weights of the reference's shapes (kaiming-uniform-like range), bf16-representable."""
    rng = np.random.RandomState(seed)
    def u(shape, fan_in):
        b = 1.0 / np.sqrt(fan_in) if scale is None else scale
        return bf16_round(rng.uniform(-b, b, size=shape).astype(np.float32))
    w = O.MoEWeights(u((e, h, ie), h), u((e, h, ie), h), u((e, ie, h), ie),
                     u((s, h), h) if s else None, u((s, h), h) if s else None, u((h, s), s) if s else None)
    cfg = O.MoEConfig(h, e, ie, v, top_k=k, primary_weight=pw if k > 1 else 1.0, shared_inter=s,
                      shared_output_scale=ss, routed_output_scale=rs)
    table = O.synthetic_hash_table(v, e, k, layer) if e > 1 else np.zeros((k, v), dtype=np.int64)
    if zipf:
        ranks = rng.zipf(1.1, size=n).astype(np.int64)
        ids = np.random.RandomState(7).permutation(v)[(ranks - 1) % v]
    else:
        ids = np.random.RandomState(1234 + seed).randint(0, v, size=n).astype(np.int64)
    x = bf16_round(np.random.RandomState(seed + 99).randn(n, h).astype(np.float32))
    return cfg, w, table, ids, x


def to_cuda(a, dtype=None):
    """Upload to the CURRENT CUDA device."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def build_gpu_block(cfg: O.MoEConfig, w: O.MoEWeights, table, mode="bf16", wq: O.MoEWeightsInt8 = None, device=None):
    import torch
    from vllm_i64 import _C
    from vllm_i64.kernels.block import I64Block
    m = _C.MODE_W8A8 if mode == "int8" else _C.MODE_BF16
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    blk = I64Block(cfg.hidden_size, cfg.num_experts, cfg.expert_inter, cfg.shared_inter, cfg.vocab_size, cfg.top_k,
                   cfg.route_weights(), cfg.routed_output_scale, cfg.shared_output_scale, m, device)
    has_sh = cfg.shared_inter > 0
    if mode == "bf16":
        blk.load_bf16(to_cuda(w.gate), to_cuda(w.up), to_cuda(w.down), *(
            (to_cuda(w.shared_gate), to_cuda(w.shared_up), to_cuda(w.shared_down)) if has_sh else (None,) * 3))
    else:
        sh = (to_cuda(wq.sh13_q), to_cuda(wq.sh13_s), to_cuda(wq.sh2_q), to_cuda(wq.sh2_s)) if has_sh else (None,) * 4
        blk.load_w8a8(to_cuda(wq.w13_q), to_cuda(wq.w13_s), to_cuda(wq.w2_q), to_cuda(wq.w2_s), *sh)
    blk.set_table(torch.from_numpy(np.asarray(table, dtype=np.int64)))
    return blk


TOL_MAX, TOL_MEAN = 1e-2, 1e-3     # north_star: ≤1e-2 max-rel, ≤1e-3 mean-rel


def assert_close(got, ref, what=""):
    mx, mean = O.rel_errors(got, ref)
    assert mx <= TOL_MAX and mean <= TOL_MEAN, f"{what}: max-rel {mx:.3e} mean-rel {mean:.3e}"
    return mx, mean
