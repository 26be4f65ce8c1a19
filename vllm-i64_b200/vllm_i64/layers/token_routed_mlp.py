"""``TokenRoutedMLP`` — drop-in for layers/token_routed_mlp.py:9-258 on the sm_100a block.

Same constructor, buffers (``token_to_expert``, ``topk_token_to_expert``), parameters (``gate_proj_w``,
``up_proj_w``, ``down_proj_w``, ``shared_gate/up/down.weight``, ``shared_output_gate``, ``routed_output_gate``) and
methods (``route``, ``expert_forward``, ``forward``, ``load_full_weights``, ``fuse_shared_gate_up``,
``materialize_expert_linears``) as the reference, so ``core/loader.py:374-492`` and the model
(``models/complexity_deep/model.py:321-343,362``) work unchanged.  The forward is ONE C-ABI call
(``i64b_moe_forward``): routing → deterministic dispatch → grouped tcgen05 GEMMs → combine, with no host sync
(the reference syncs 2·E·K times per layer, token_routed_mlp.py:136-151).  CPU tensors raise: this build has no
CPU fallback.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from vllm_i64 import _C
from vllm_i64.kernels import torch_ops
from vllm_i64.kernels.block import I64Block
from vllm_i64.parallel.tensor_parallel import all_reduce, get_tp


class TokenRoutedMLP(torch_ops.TracedLayer, nn.Module):
    def __init__(self, hidden_size: int, intermediate_size: int, num_experts: int, vocab_size: int,
                 shared_expert: bool = False, shared_intermediate_size: int = 0, top_k: int = 1,
                 top_k_primary_weight: Optional[float] = None, use_shared_routed_gates: bool = False,
                 shared_gate_init: float = 1.0, routed_gate_init: float = 1.0, shared_output_scale: float = 1.0,
                 routed_output_scale: float = 1.0):
        super().__init__()
        tp = get_tp()
        self.hidden_size = hidden_size
        self.num_experts = num_experts
        self.vocab_size = vocab_size
        self.tp_size = tp.tp_size
        self.tp_rank = tp.tp_rank
        self.top_k = max(1, int(top_k))
        if self.top_k > self.num_experts:
            raise ValueError(f"top_k ({self.top_k}) cannot exceed num_experts ({self.num_experts})")
        if top_k_primary_weight is None:
            top_k_primary_weight = 0.95
        self.primary_weight = (min(1.0, max(0.0, float(top_k_primary_weight))) if self.top_k > 1 else 1.0)
        self.shared_output_scale = float(shared_output_scale)
        self.routed_output_scale = float(routed_output_scale)
        self.full_expert_inter = intermediate_size // num_experts
        self.expert_inter = self.full_expert_inter // tp.tp_size

        self.gate_proj_w = nn.Parameter(torch.empty(num_experts, hidden_size, self.expert_inter))
        self.up_proj_w = nn.Parameter(torch.empty(num_experts, hidden_size, self.expert_inter))
        self.down_proj_w = nn.Parameter(torch.empty(num_experts, self.expert_inter, hidden_size))

        self.use_shared_expert = shared_expert
        self.use_shared_routed_gates = bool(use_shared_routed_gates)
        self.shared_size = 0
        if shared_expert:
            self.shared_size = shared_intermediate_size if shared_intermediate_size > 0 else self.full_expert_inter
            self.shared_gate = nn.Linear(hidden_size, self.shared_size, bias=False)
            self.shared_up = nn.Linear(hidden_size, self.shared_size, bias=False)
            self.shared_down = nn.Linear(self.shared_size, hidden_size, bias=False)
            if self.use_shared_routed_gates:
                self.shared_output_gate = nn.Parameter(torch.tensor(float(shared_gate_init)))
                self.routed_output_gate = nn.Parameter(torch.tensor(float(routed_gate_init)))

        ar = torch.arange(vocab_size, dtype=torch.long)
        self.register_buffer("token_to_expert", ar % num_experts)
        self.register_buffer("topk_token_to_expert", torch.stack([(ar + r) % num_experts for r in range(self.top_k)]))

        for w in (self.gate_proj_w, self.up_proj_w, self.down_proj_w):
            nn.init.kaiming_uniform_(w, a=5 ** 0.5)

        self.quantization = "none"        # "none" (bf16) | "int8" (dynamic W8A8)
        torch_ops.register_layer(self)
        self._block: Optional[I64Block] = None
        self._block_key = None

    # ------------------------------------------------------------------ weights in → packed block
    def route_weights(self):
        if self.top_k == 1:
            return [1.0]
        sec = (1.0 - self.primary_weight) / (self.top_k - 1)
        return [self.primary_weight] + [sec] * (self.top_k - 1)

    def _expert_tensors(self):
        """(gate [E,H,I], up [E,H,I], down [E,I,H]) whichever form the module currently holds."""
        if hasattr(self, "expert_gate_up"):
            i = self.expert_inter
            gu = torch.stack([m.weight for m in self.expert_gate_up])            # [E, 2I, H]
            dn = torch.stack([m.weight for m in self.expert_down])               # [E, H, I]
            return gu[:, :i].transpose(1, 2), gu[:, i:].transpose(1, 2), dn.transpose(1, 2)
        if hasattr(self, "gate_up_proj_w"):
            i = self.expert_inter
            return self.gate_up_proj_w[..., :i], self.gate_up_proj_w[..., i:], self.down_proj_w
        return self.gate_proj_w, self.up_proj_w, self.down_proj_w

    def _shared_tensors(self):
        if not self.use_shared_expert:
            return None, None, None
        if hasattr(self, "shared_gate_up"):
            s = self.shared_size
            return self.shared_gate_up.weight[:s], self.shared_gate_up.weight[s:], self.shared_down.weight
        return self.shared_gate.weight, self.shared_up.weight, self.shared_down.weight

    def _scales(self):
        if self.use_shared_expert and self.use_shared_routed_gates:
            # one D2H read per (re)pack, never per forward: the learned gates are constants at inference time
            g = torch.stack([self.routed_output_gate.detach().float(), self.shared_output_gate.detach().float()]).tolist()
            return g[0], g[1]
        return self.routed_output_scale, (self.shared_output_scale if self.use_shared_expert else 1.0)

    def _param_tensors(self):
        """The live parameter / buffer tensors the packed block is derived from, WITHOUT building temporaries (the
        materialised form holds one nn.Linear per expert)."""
        ts = []
        if hasattr(self, "expert_gate_up"):
            ts += [m.weight for m in self.expert_gate_up] + [m.weight for m in self.expert_down]
        elif hasattr(self, "gate_up_proj_w"):
            ts += [self.gate_up_proj_w, self.down_proj_w]
        else:
            ts += [self.gate_proj_w, self.up_proj_w, self.down_proj_w]
        if self.use_shared_expert:
            ts += ([self.shared_gate_up.weight] if hasattr(self, "shared_gate_up")
                   else [self.shared_gate.weight, self.shared_up.weight]) + [self.shared_down.weight]
            if self.use_shared_routed_gates:
                ts += [self.routed_output_gate, self.shared_output_gate]
        ts.append(self.topk_token_to_expert)
        return ts

    def _state_key(self):
        """Identity + storage + in-place version of every source tensor.  Catches ``param.copy_()``, ``load_state_dict``,
        re-assignment and ``.to(device)``.  It cannot see writes through ``.data`` (``p.data.copy_(w)`` does not bump
        ``_version``): the module's own loaders (``load_full_weights``, ``_load_from_state_dict``) invalidate explicitly,
        anything else that rewrites weights in place after the first forward must call ``invalidate()``."""
        return (self.quantization, self.tp_size,
                tuple((id(t), t.data_ptr(), t._version, tuple(t.shape)) for t in self._param_tensors()))

    def invalidate(self) -> None:
        """Drop the packed block: the next forward re-packs from the current parameters."""
        self._block = None
        self._block_key = None

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        self.invalidate()

    def _ensure_block(self, device: torch.device) -> I64Block:
        key = (device, self._state_key())
        if self._block is not None and self._block_key == key:
            return self._block
        rs, ss = self._scales()
        if self.tp_size > 1:
            ss = ss / self.tp_size     # the shared expert is replicated: its term is summed tp_size times below
        mode = {"int8": _C.MODE_W8A8, "fp8": _C.MODE_FP8}.get(self.quantization, _C.MODE_BF16)
        blk = I64Block(self.hidden_size, self.num_experts, self.expert_inter, self.shared_size, self.vocab_size,
                       self.top_k, self.route_weights(), rs, ss, mode, device)
        gate, up, down = self._expert_tensors()
        sg, su, sd = self._shared_tensors()
        with torch.no_grad():
            if mode == _C.MODE_BF16:
                blk.load_bf16(gate, up, down, sg, su, sd)
            else:
                if mode == _C.MODE_FP8:
                    from vllm_i64.core.fp8 import quantize_fp8 as quantize_int8      # same call shape: rows -> (q, scale)
                else:
                    from vllm_i64.core.quantization import quantize_int8
                # materialised (out,in) forms, layers/token_routed_mlp.py:232-242 of the reference
                w13 = torch.cat([gate.transpose(1, 2), up.transpose(1, 2)], dim=1).float().contiguous()
                w2 = down.transpose(1, 2).float().contiguous()
                w13_q, w13_s = quantize_int8(w13)
                w2_q, w2_s = quantize_int8(w2)
                sh = [None] * 4
                if sg is not None:
                    sh13_q, sh13_s = quantize_int8(torch.cat([sg, su], dim=0).float().contiguous())
                    sh2_q, sh2_s = quantize_int8(sd.float().contiguous())
                    sh = [sh13_q, sh13_s, sh2_q, sh2_s]
                (blk.load_fp8 if mode == _C.MODE_FP8 else blk.load_w8a8)(w13_q, w13_s, w2_q, w2_s, *sh)
            blk.set_table(self.topk_token_to_expert)
        self._block, self._block_key = blk, key
        return blk

    def quantize_fp8(self) -> "TokenRoutedMLP":
        """Switch the block to FP8: E4M3 weights per out-row, activations quantised per token, fp32 accumulation — the recipe
        of core/fp8.py + layers/dense_mlp.py:75-83 (fused gate|up, SiLU·up, per-row re-quantisation, down) applied to every
        expert's materialised linears and to the shared expert."""
        self.quantization = "fp8"
        self._block = None
        return self

    def quantize_w8a8(self) -> "TokenRoutedMLP":
        """Switch the block to dynamic W8A8 (per-token activation scale × per-out-channel weight scale, INT32
        accumulate, Q7-LUT SiLU) — the semantics of core/quantization.py:80-286 + layers/moe.py:114-122."""
        self.quantization = "int8"
        self._block = None
        return self

    # ------------------------------------------------------------------ reference API
    def route(self, token_ids: Optional[torch.Tensor], num_tokens: int, device: torch.device) -> torch.Tensor:
        """``[top_k, num_tokens]`` int64 expert lookup (layers/token_routed_mlp.py:113-130)."""
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("TokenRoutedMLP.route: CUDA only in this build (no CPU fallback)")
        blk = self._ensure_block(device)
        out = torch.empty(self.top_k, num_tokens, dtype=torch.int64, device=device)
        ids = None
        if token_ids is not None:
            ids = token_ids.to(device=device, dtype=torch.int64).contiguous()
        with torch.cuda.device(device):
            _C.check(_C.lib().i64b_route_table(_C.ptr(ids), blk._table_u8.data_ptr(), out.data_ptr(), num_tokens,
                                               self.vocab_size, self.top_k, self.num_experts, _C.stream_ptr(device)))
        return out

    def expert_forward(self, x: torch.Tensor, expert_ids: torch.Tensor) -> torch.Tensor:
        """Routed experts only, weight 1 (layers/token_routed_mlp.py:132-152) + the TP all-reduce."""
        _C.require_cuda(x, expert_ids)
        blk = self._ensure_block(x.device)
        x2 = x.reshape(-1, self.hidden_size)
        odt = torch.bfloat16 if (x.dtype == torch.bfloat16 and self.tp_size == 1) else torch.float32
        out = blk.forward_routed(x2, expert_ids.reshape(1, -1), [1.0], 1.0, 1.0, False, odt)
        out = all_reduce(out)
        return out.to(x.dtype).reshape(x.shape)

    def forward(self, x: torch.Tensor, token_ids: Optional[torch.Tensor] = None, x_preq=None, **kwargs) -> torch.Tensor:
        """``x_preq = (int8 [N,H], scale)`` (optional, W8A8 mode): rows already quantised by the fused RMSNorm+quant in
        front of the block (``get_i64_cuda_ops().add_rmsnorm(..., quantize=True)``); the block then skips its own
        absmax / quantise pass.  Same result bit for bit.  While ``torch.compile`` traces, the whole forward is one
        ``vllm_i64::token_routed_mlp`` node (kernels/torch_ops.py; fused_i64_moe.py:123-142 of the reference)."""
        if torch.compiler.is_compiling():
            return torch_ops.token_routed_mlp(x, token_ids, x_preq, self)
        return self._forward_eager(x, token_ids, x_preq)

    def _forward_eager(self, x: torch.Tensor, token_ids: Optional[torch.Tensor] = None, x_preq=None) -> torch.Tensor:
        _C.require_cuda(x)
        blk = self._ensure_block(x.device)
        x2 = x.reshape(-1, self.hidden_size)
        ids = None if token_ids is None else token_ids.reshape(-1)
        use_preq = x_preq is not None and self.quantization == "int8"

        def fwd(odt):
            if use_preq:
                return blk.forward_preq(x_preq[0].reshape(x2.shape), x_preq[1], ids, odt)
            return blk.forward(x2, ids, odt)
        if self.tp_size > 1:
            out = all_reduce(fwd(torch.float32))
            return out.to(x.dtype).reshape(x.shape)
        odt = torch.bfloat16 if x.dtype == torch.bfloat16 else torch.float32
        out = fwd(odt)
        return (out if out.dtype == x.dtype else out.to(x.dtype)).reshape(x.shape)

    def fuse_shared_gate_up(self) -> bool:
        """Fuse the shared SwiGLU input projections (layers/token_routed_mlp.py:189-208).  The packed block always
        holds gate|up fused; this only changes the module's parameter form, as in the reference."""
        if not self.use_shared_expert or hasattr(self, "shared_gate_up"):
            return False
        fused = nn.Linear(self.shared_gate.in_features, self.shared_gate.out_features + self.shared_up.out_features,
                          bias=False, device=self.shared_gate.weight.device, dtype=self.shared_gate.weight.dtype)
        with torch.no_grad():
            fused.weight.copy_(torch.cat([self.shared_gate.weight, self.shared_up.weight], dim=0))
        self.shared_gate_up = fused
        del self.shared_gate
        del self.shared_up
        self._block = None
        return True

    def materialize_expert_linears(self) -> bool:
        """Per-expert ``nn.Linear`` modules ``(2I,H)`` / ``(H,I)`` (layers/token_routed_mlp.py:210-250)."""
        if hasattr(self, "expert_gate_up"):
            return False
        gu, dn = [], []
        for e in range(self.num_experts):
            a = nn.Linear(self.hidden_size, 2 * self.expert_inter, bias=False, device=self.gate_proj_w.device,
                          dtype=self.gate_proj_w.dtype)
            b = nn.Linear(self.expert_inter, self.hidden_size, bias=False, device=self.down_proj_w.device,
                          dtype=self.down_proj_w.dtype)
            with torch.no_grad():
                a.weight.copy_(torch.cat([self.gate_proj_w[e].T, self.up_proj_w[e].T], dim=0))
                b.weight.copy_(self.down_proj_w[e].T)
            gu.append(a)
            dn.append(b)
        self.expert_gate_up = nn.ModuleList(gu)
        self.expert_down = nn.ModuleList(dn)
        del self.gate_proj_w
        del self.up_proj_w
        del self.down_proj_w
        self._block = None
        return True

    def load_full_weights(self, full_gate: torch.Tensor, full_up: torch.Tensor, full_down: torch.Tensor) -> None:
        """Unsharded expert weights → this TP rank's slice (layers/token_routed_mlp.py:252-258)."""
        lo = self.tp_rank * self.expert_inter
        hi = lo + self.expert_inter
        self.gate_proj_w.data.copy_(full_gate[:, :, lo:hi])
        self.up_proj_w.data.copy_(full_up[:, :, lo:hi])
        self.down_proj_w.data.copy_(full_down[:, lo:hi, :])
        self.invalidate()

    # per-expert token histogram of the last forward (what /v1/experts re-derives on the host, api/_admin.py:464-562)
    def last_expert_counts(self) -> torch.Tensor:
        if self._block is None:
            raise RuntimeError("no forward has run yet")
        return self._block.tap(_C.TAP_COUNTS, (self.top_k, self.num_experts), torch.int32)
