"""``DenseMLP`` / ``DenseSwiGLUMLP`` — drop-ins for layers/dense_mlp.py:33-138 and
models/complexity_deep/model.py:243-276 (the FFN ``dense-306`` runs; BASELINE config 4).

``down(silu(gate(x)) * up(x))`` as two grouped-GEMM launches of the same tcgen05 kernels the MoE block uses
(one "expert" holding every row): GEMM1 with the SwiGLU epilogue, GEMM2.  ``quantize_w8a8()`` selects the
reference's ``_forward_int8_fused`` semantics (one activation quant → fused gate|up INT8 GEMM → Q7-LUT SiLU·up →
re-quant → down INT8 GEMM).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from vllm_i64 import _C
from vllm_i64.kernels import torch_ops
from vllm_i64.kernels.block import I64Block
from vllm_i64.parallel.tensor_parallel import ColumnParallelLinear, RowParallelLinear, all_reduce, get_tp


class _DenseBlockMixin(torch_ops.TracedLayer):
    quantization = "none"
    _block: Optional[I64Block] = None
    _block_key = None

    def _weights(self):
        raise NotImplementedError

    def _ensure_block(self, device: torch.device) -> I64Block:
        gate, up, down = self._weights()
        # identity + storage + in-place version; writes through ``.data`` are invisible to it: see invalidate()
        ref_fp8 = getattr(self, "gate_fp8", None) is not None      # buffers left by the reference's _quantize_dense_mlp_fp8
        ref_int8 = getattr(self, "gate_up_int8", None) is not None  # ... and by _quantize_dense_mlp(model, "int8")
        if ref_fp8:
            self.quantization = "fp8"
        elif ref_int8:
            self.quantization = "int8"
        if getattr(self, "gate_int4", None) is not None:
            # buffers of _quantize_dense_mlp(model, "int4") (core/loader.py:685-697): there is no INT4 kernel in this build, and
            # running the float weights instead would silently ignore the quantisation the user asked for
            raise NotImplementedError(
                "vllm_i64 (B200 build): INT4 dense MLP buffers (gate_int4 ...) are present but this build has no INT4 "
                "kernel; load with --quantization int8 or fp8, or without quantisation")
        # the packed block is made from the reference loader's quantised buffers when they exist: they are part of the key, so
        # re-registering or rewriting them repacks
        qnames = (("gate_fp8", "gate_fp8_scale", "up_fp8", "up_fp8_scale", "down_fp8", "down_fp8_scale") if ref_fp8 else
                  ("gate_up_int8", "gate_up_scale", "down_int8", "down_scale") if ref_int8 else ())
        qbufs = tuple(getattr(self, n, None) for n in qnames)
        key = (device, self.quantization, tuple((t.data_ptr(), t._version, tuple(t.shape)) for t in (gate, up, down)),
               tuple(id(p) for p in self.parameters()),
               tuple((n, b.data_ptr(), b._version, tuple(b.shape)) for n, b in zip(qnames, qbufs) if b is not None))
        if self._block is not None and self._block_key == key:
            return self._block
        inter, hidden = gate.shape
        mode = {"int8": _C.MODE_W8A8, "fp8": _C.MODE_FP8}.get(self.quantization, _C.MODE_BF16)
        blk = I64Block(hidden, 0, 0, inter, 1, 0, [], 1.0, 1.0, mode, device)
        with torch.no_grad():
            if mode == _C.MODE_BF16:
                blk.load_bf16(None, None, None, gate, up, down)
            elif mode == _C.MODE_FP8:
                if ref_fp8:         # core/loader.py:958-987: the bytes and scales the reference loader made, as they are
                    q13 = torch.cat([self.gate_fp8.view(torch.uint8), self.up_fp8.view(torch.uint8)], dim=0).to(device)
                    s13 = torch.cat([self.gate_fp8_scale, self.up_fp8_scale], dim=0).to(device)
                    q2, s2 = self.down_fp8.view(torch.uint8).to(device), self.down_fp8_scale.to(device)
                else:
                    from vllm_i64.core.fp8 import quantize_fp8
                    q13, s13 = quantize_fp8(torch.cat([gate, up], dim=0).float().contiguous())
                    q2, s2 = quantize_fp8(down.float().contiguous())
                blk.load_fp8(None, None, None, None, q13, s13, q2, s2)
            else:
                if ref_int8:        # core/loader.py:670-683: fused gate|up int8 rows + scales, down int8 + scales
                    q13, s13 = self.gate_up_int8.to(device), self.gate_up_scale.to(device).float()
                    q2, s2 = self.down_int8.to(device), self.down_scale.to(device).float()
                else:
                    from vllm_i64.core.quantization import quantize_int8
                    q13, s13 = quantize_int8(torch.cat([gate, up], dim=0).float().contiguous())
                    q2, s2 = quantize_int8(down.float().contiguous())
                blk.load_w8a8(None, None, None, None, q13, s13, q2, s2)
        self._block, self._block_key = blk, key
        return blk

    def _forward_eager(self, x, token_ids=None, x_preq=None):
        return self._run(x, x_preq)

    def quantize_w8a8(self):
        self.quantization = "int8"
        self._block = None
        return self

    def quantize_fp8(self):
        """layers/dense_mlp.py:75-83 (``_forward_fp8``) as one FP8 block: fused gate|up GEMM, SiLU·up, per-row re-quantisation,
        down GEMM on tcgen05 ``kind::f8f6f4``."""
        self.quantization = "fp8"
        self._block = None
        return self

    def invalidate(self) -> None:
        """Drop the packed block (call after rewriting weights through ``.data``; ``load_state_dict`` does it itself)."""
        self._block = None
        self._block_key = None

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        self.invalidate()

    def _run(self, x: torch.Tensor, x_preq=None) -> torch.Tensor:
        _C.require_cuda(x)
        blk = self._ensure_block(x.device)
        x2 = x.reshape(-1, x.shape[-1])
        tp = get_tp()
        use_preq = x_preq is not None and self.quantization == "int8"

        def fwd(odt):
            if use_preq:          # (int8 rows, per-token scale) from the fused RMSNorm+quant: no second quantisation pass
                return blk.forward_preq(x_preq[0].reshape(x2.shape), x_preq[1], None, odt)
            return blk.forward(x2, None, odt)
        if tp.tp_size > 1:
            out = all_reduce(fwd(torch.float32))
            return out.to(x.dtype).reshape(x.shape)
        odt = torch.bfloat16 if x.dtype == torch.bfloat16 else torch.float32
        out = fwd(odt)
        return (out if out.dtype == x.dtype else out.to(x.dtype)).reshape(x.shape)


class DenseMLP(_DenseBlockMixin, nn.Module):
    """TP-aware dense SwiGLU (layers/dense_mlp.py:33-138): gate/up column-parallel, down row-parallel."""

    def __init__(self, hidden_size: int, intermediate_size: int):
        super().__init__()
        torch_ops.register_layer(self)
        tp = get_tp()
        self.hidden_size = hidden_size
        self.intermediate_size = intermediate_size
        self.inter_per_tp = intermediate_size // tp.tp_size
        self.gate_proj = ColumnParallelLinear(hidden_size, intermediate_size, bias=False)
        self.up_proj = ColumnParallelLinear(hidden_size, intermediate_size, bias=False)
        self.down_proj = RowParallelLinear(intermediate_size, hidden_size, bias=False)

    def _weights(self):
        return self.gate_proj.linear.weight, self.up_proj.linear.weight, self.down_proj.linear.weight

    def forward(self, x: torch.Tensor, x_preq=None, **kwargs) -> torch.Tensor:
        """``x_preq = (int8 [N,H], scale [N,1])`` from the fused RMSNorm+quant (layers/dense_mlp.py:46-58 of the
        reference): in W8A8 mode the block consumes it directly (``i64b_moe_forward_preq``) and skips its own
        quantisation pass; the result is bit-identical either way (same quantisation rule)."""
        if torch.compiler.is_compiling():
            return torch_ops.token_routed_mlp(x, None, x_preq, self)
        return self._run(x, x_preq)


class DenseSwiGLUMLP(_DenseBlockMixin, nn.Module):
    """models/complexity_deep/model.py:243-276 (what ``dense-306`` instantiates)."""

    def __init__(self, config):
        super().__init__()
        torch_ops.register_layer(self)
        self.gate_proj = nn.Linear(config.hidden_size, config.intermediate_size, bias=False)
        self.up_proj = nn.Linear(config.hidden_size, config.intermediate_size, bias=False)
        self.down_proj = nn.Linear(config.intermediate_size, config.hidden_size, bias=False)

    def _weights(self):
        if hasattr(self, "gate_up_proj"):
            i = self.gate_up_proj.out_features // 2
            return self.gate_up_proj.weight[:i], self.gate_up_proj.weight[i:], self.down_proj.weight
        return self.gate_proj.weight, self.up_proj.weight, self.down_proj.weight

    def forward(self, x, token_ids=None, **kwargs):
        if torch.compiler.is_compiling():
            return torch_ops.token_routed_mlp(x, None, None, self)
        return self._run(x)

    def fuse_gate_up(self) -> bool:
        if hasattr(self, "gate_up_proj"):
            return False
        fused = nn.Linear(self.gate_proj.in_features, self.gate_proj.out_features + self.up_proj.out_features,
                          bias=False, device=self.gate_proj.weight.device, dtype=self.gate_proj.weight.dtype)
        with torch.no_grad():
            fused.weight.copy_(torch.cat([self.gate_proj.weight, self.up_proj.weight], dim=0))
        self.gate_up_proj = fused
        del self.gate_proj
        del self.up_proj
        self._block = None
        return True
