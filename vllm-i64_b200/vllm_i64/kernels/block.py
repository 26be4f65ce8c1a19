"""``I64Block`` — Python owner of one ``i64b_layer`` handle (include/i64b.h): packed weights, routing table,
workspace and the forward calls.  All device memory is torch-owned; the C side never allocates."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from vllm_i64 import _C


class I64Block:
    def __init__(self, hidden_size: int, num_experts: int, expert_inter: int, shared_inter: int, vocab_size: int,
                 top_k: int, route_weights: Sequence[float], routed_scale: float, shared_scale: float,
                 mode: int, device: torch.device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("I64Block needs a CUDA device: this build has no CPU path")
        cfg = _C.MoEConfig()
        cfg.hidden_size, cfg.num_experts, cfg.expert_inter = int(hidden_size), int(num_experts), int(expert_inter)
        cfg.shared_inter, cfg.vocab_size, cfg.top_k, cfg.mode = int(shared_inter), int(vocab_size), int(top_k), int(mode)
        for i in range(4):
            cfg.route_weight[i] = float(route_weights[i]) if i < len(route_weights) else 0.0
        cfg.routed_scale, cfg.shared_scale = float(routed_scale), float(shared_scale)
        self.cfg = cfg
        self.mode = int(mode)
        self.top_k = int(top_k) if num_experts > 0 else 0
        self._h = C.c_void_p()
        _C.check(_C.lib().i64b_moe_create(C.byref(cfg), C.byref(self._h)))
        self._packed: Optional[torch.Tensor] = None
        self._table_u8: Optional[torch.Tensor] = None

    def __deepcopy__(self, memo):
        """A packed block is not copied with its module (the handle owns device memory): the copy packs its own lazily."""
        return None

    def __reduce__(self):
        return (type(None), ())

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _C.lib().i64b_moe_destroy(h)
            except Exception:
                pass
            self._h = C.c_void_p()

    # ---- parameters
    def _alloc_packed(self):
        nbytes = _C.lib().i64b_moe_packed_bytes(self._h)
        self._packed = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)

    @staticmethod
    def _bf16(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        return None if t is None else t.detach().to(torch.bfloat16).contiguous()

    def load_bf16(self, gate, up, down, shared_gate=None, shared_up=None, shared_down=None) -> None:
        """gate/up ``[E,H,I]``, down ``[E,I,H]``, shared nn.Linear weights ``(out,in)`` (reference layouts)."""
        with torch.cuda.device(self.device):
            self._alloc_packed()
            ts = [self._bf16(t) for t in (gate, up, down, shared_gate, shared_up, shared_down)]
            _C.check(_C.lib().i64b_moe_load_bf16(self._h, *[_C.ptr(t) for t in ts], self._packed.data_ptr(),
                                                 _C.stream_ptr(self.device)))
            self._keep = ts    # inputs must outlive the stream-ordered repack

    def load_w8a8(self, w13_q, w13_s, w2_q, w2_s, sh13_q=None, sh13_s=None, sh2_q=None, sh2_s=None) -> None:
        """Materialised ``(out,in)`` int8 weights + per-out-row fp32 scales (see include/i64b.h)."""
        with torch.cuda.device(self.device):
            self._alloc_packed()
            ts = [None if t is None else t.detach().contiguous() for t in
                  (w13_q, w13_s, w2_q, w2_s, sh13_q, sh13_s, sh2_q, sh2_s)]
            _C.check(_C.lib().i64b_moe_load_w8a8(self._h, *[_C.ptr(t) for t in ts], self._packed.data_ptr(),
                                                 _C.stream_ptr(self.device)))
            self._keep = ts

    def load_fp8(self, w13_q, w13_s, w2_q, w2_s, sh13_q=None, sh13_s=None, sh2_q=None, sh2_s=None) -> None:
        """Materialised ``(out,in)`` E4M3 weights (``torch.float8_e4m3fn`` or their uint8 view) + per-out-row fp32 scales
        (``quantize_fp8``, core/fp8.py:75-110); same layouts as :meth:`load_w8a8`."""
        def as_u8(t):
            if t is None:
                return None
            t = t.detach().contiguous()
            return t.view(torch.uint8) if t.dtype != torch.uint8 else t
        with torch.cuda.device(self.device):
            self._alloc_packed()
            ts = [as_u8(w13_q), None if w13_s is None else w13_s.detach().float().contiguous(),
                  as_u8(w2_q), None if w2_s is None else w2_s.detach().float().contiguous(),
                  as_u8(sh13_q), None if sh13_s is None else sh13_s.detach().float().contiguous(),
                  as_u8(sh2_q), None if sh2_s is None else sh2_s.detach().float().contiguous()]
            _C.check(_C.lib().i64b_moe_load_fp8(self._h, *[_C.ptr(t) for t in ts], self._packed.data_ptr(),
                                                _C.stream_ptr(self.device)))
            self._keep = ts

    def set_table(self, table_i64: torch.Tensor) -> None:
        """``topk_token_to_expert [K,V]`` int64 → uint8 device table."""
        t = table_i64.detach().to(self.device, torch.int64).contiguous()
        k, v = t.shape
        out = torch.empty(k, v, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _C.check(_C.lib().i64b_pack_table(t.data_ptr(), out.data_ptr(), k, v, max(1, self.cfg.num_experts),
                                              _C.stream_ptr(self.device)))
        self._table_u8 = out
        self._table_src = t
        _C.check(_C.lib().i64b_moe_set_table(self._h, out.data_ptr()))

    def set_scales(self, route_weights: Sequence[float], routed_scale: float, shared_scale: float) -> None:
        arr = (C.c_float * 4)(*[float(route_weights[i]) if i < len(route_weights) else 0.0 for i in range(4)])
        _C.check(_C.lib().i64b_moe_set_scales(self._h, arr, self.top_k, float(routed_scale), float(shared_scale)))

    # ---- forward
    def _out(self, n: int, out_dtype: torch.dtype) -> torch.Tensor:
        if out_dtype not in (torch.bfloat16, torch.float32):
            raise TypeError("output dtype must be bfloat16 or float32")
        return torch.empty(n, self.cfg.hidden_size, dtype=out_dtype, device=self.device)

    def forward(self, x: torch.Tensor, token_ids: Optional[torch.Tensor], out_dtype=torch.bfloat16,
                out: Optional[torch.Tensor] = None, stages: int = _C.STAGE_ALL) -> torch.Tensor:
        _C.require_cuda(x, token_ids)
        if x.dtype != torch.bfloat16 or not x.is_contiguous():
            x = x.to(torch.bfloat16).contiguous()
        n = x.shape[0]
        if token_ids is not None:
            if token_ids.dtype != torch.int64 or not token_ids.is_contiguous():
                token_ids = token_ids.to(torch.int64).contiguous()
            if token_ids.numel() != n:
                raise RuntimeError("token_ids must have one id per row of x")
        if out is None:
            out = self._out(n, out_dtype)
        lib = _C.lib()
        with torch.cuda.device(self.device):
            ws = _C.workspace(self.device, lib.i64b_moe_workspace_bytes(self._h, n))
            _C.check(lib.i64b_moe_forward_stages(self._h, x.data_ptr(), _C.ptr(token_ids), out.data_ptr(),
                                                 _C.dtype_code(out), n, ws.data_ptr(), ws.numel(), int(stages),
                                                 _C.stream_ptr(self.device)))
        return out

    def forward_preq(self, x_q: torch.Tensor, x_scale: torch.Tensor, token_ids: Optional[torch.Tensor],
                     out_dtype=torch.bfloat16, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """W8A8 forward on rows the caller has already quantised (``x_preq = (int8 [N,H], fp32 scale [N])`` from the
        fused RMSNorm+quant): no absmax / quantise pass, bit-identical to ``forward`` on the un-quantised rows."""
        _C.require_cuda(x_q, x_scale, token_ids)
        if x_q.dtype != torch.int8:
            raise TypeError("x_q must be int8")
        x_q = x_q.contiguous()
        x_scale = x_scale.reshape(-1).to(torch.float32).contiguous()
        n = x_q.shape[0]
        if x_scale.numel() != n:
            raise RuntimeError("x_scale must hold one scale per row of x_q")
        if token_ids is not None:
            token_ids = token_ids.to(torch.int64).contiguous()
        if out is None:
            out = self._out(n, out_dtype)
        lib = _C.lib()
        with torch.cuda.device(self.device):
            ws = _C.workspace(self.device, lib.i64b_moe_workspace_bytes(self._h, n))
            _C.check(lib.i64b_moe_forward_preq(self._h, x_q.data_ptr(), x_scale.data_ptr(), _C.ptr(token_ids),
                                               out.data_ptr(), _C.dtype_code(out), n, ws.data_ptr(), ws.numel(),
                                               _C.stream_ptr(self.device)))
        return out

    def uses_fused(self, n: int) -> bool:
        return bool(_C.lib().i64b_moe_uses_fused(self._h, int(n)))

    def fuses_combine(self, n: int) -> bool:
        return bool(_C.lib().i64b_moe_fuses_combine(self._h, int(n)))

    def tri_form(self, n: int) -> int:
        """0: GEMM2 + combine; 1: fused down projection, three accumulators; 2: one accumulator (power-of-two scales)."""
        return int(_C.lib().i64b_moe_tri_form(self._h, int(n)))

    def forward_routed(self, x: torch.Tensor, routes: torch.Tensor, route_weights: Sequence[float],
                       routed_scale: float = 1.0, shared_scale: float = 1.0, use_shared: bool = False,
                       out_dtype=torch.bfloat16) -> torch.Tensor:
        """Caller-supplied routes ``[K,N]`` (any integer dtype)."""
        _C.require_cuda(x, routes)
        if x.dtype != torch.bfloat16 or not x.is_contiguous():
            x = x.to(torch.bfloat16).contiguous()
        r = routes.to(torch.int32).contiguous()
        if r.dim() == 1:
            r = r.view(1, -1)
        k, n = r.shape
        if n != x.shape[0]:
            raise RuntimeError("routes must be [K, N] with N == x.shape[0]")
        out = self._out(n, out_dtype)
        w = (C.c_float * 4)(*[float(route_weights[i]) if i < len(route_weights) else 0.0 for i in range(4)])
        lib = _C.lib()
        with torch.cuda.device(self.device):
            ws = _C.workspace(self.device, lib.i64b_moe_workspace_bytes_routed(self._h, n, k))
            _C.check(lib.i64b_moe_forward_routed(self._h, x.data_ptr(), r.data_ptr(), k, w, float(routed_scale),
                                                 float(shared_scale), int(bool(use_shared)), out.data_ptr(),
                                                 _C.dtype_code(out), n, ws.data_ptr(), ws.numel(),
                                                 _C.stream_ptr(self.device)))
        return out

    def forward_host(self, x_host: torch.Tensor, ids_host: Optional[torch.Tensor], out_host: torch.Tensor,
                     staging: dict) -> None:
        """Host-buffer call: H2D + forward + D2H enqueued on the current stream (pinned tensors recommended)."""
        n = x_host.shape[0]
        lib = _C.lib()
        with torch.cuda.device(self.device):
            ws = _C.workspace(self.device, lib.i64b_moe_workspace_bytes(self._h, n))
            _C.check(lib.i64b_moe_forward_host(
                self._h, x_host.data_ptr(), _C.ptr(ids_host), out_host.data_ptr(), _C.dtype_code(out_host), n,
                staging["x"].data_ptr(), staging["ids"].data_ptr(), staging["out"].data_ptr(), ws.data_ptr(),
                ws.numel(), _C.stream_ptr(self.device)))

    def make_staging(self, n: int, out_dtype=torch.bfloat16) -> dict:
        return {"x": torch.empty(n, self.cfg.hidden_size, dtype=torch.bfloat16, device=self.device),
                "ids": torch.empty(n, dtype=torch.int64, device=self.device),
                "out": torch.empty(n, self.cfg.hidden_size, dtype=out_dtype, device=self.device)}

    def set_next(self, nxt: Optional["I64Block"]) -> None:
        """Name the block the engine runs after this one (decoder order, reference models/complexity_deep/model.py:478-486):
        the fused decode kernel prefetches its weights into L2.  ``None`` turns the prefetch off."""
        _C.check(_C.lib().i64b_moe_set_next(self._h, nxt._h if nxt is not None else None))
        self._next = nxt          # keeps the packed weights alive while they can be prefetched

    def set_debug(self, flags: int) -> None:
        """``_C.DEBUG_KEEP_TAPS``: keep the shared expert's own rows in TAP_Y (the fused combine does not need them)."""
        _C.check(_C.lib().i64b_moe_set_debug(self._h, int(flags)))

    def set_cta_limit(self, max_ctas: int) -> None:
        """GEMM launches of this block use at most ``max_ctas`` CTAs (0 = one per SM): include/i64b.h,
        i64b_moe_set_cta_limit."""
        _C.check(_C.lib().i64b_moe_set_cta_limit(self._h, int(max_ctas)))

    def tap(self, what: int, shape, dtype: torch.dtype) -> torch.Tensor:
        with torch.cuda.device(self.device):
            return _C.tap_tensor(self._h, what, shape, dtype, self.device)


class HostStepPipeline:
    """An engine step over HOST buffers (the reference's engine hands every step's hidden states to the layer stack,
    engine/i64_engine.py; here the buffers are pinned host tensors): the step's tokens are cut into ``chunks``
    micro-batches and the host->device copy of micro-batch c+1 and the device->host copy of c-1 run on their own streams
    under the block forwards of c.  The tokens of an FFN block are independent, so the cut changes no bit of the result;
    it only keeps PCIe and the SMs busy at the same time (a 64K-token bf16 step moves 134 MB each way: 2 x 2.6 ms of
    copies around 3.5 ms of kernels when done one after the other).

    ``layers`` are applied to the SAME input rows, the last layer's output is returned (the shape of bench.py's synthetic
    step); ``chain=True`` feeds every layer the previous layer's output instead (a decoder stack without attention).
    No host synchronisation; capturable in a CUDA graph.

    ``chunks`` is a count (equal micro-batches) or a sequence of relative sizes, e.g. ``(1, 2, 2, 2, 1)``: the first
    micro-batch's copy-in and the last one's copy-out are the only copies nothing hides, so small micro-batches at both
    ends and large (more efficient) ones in the middle shorten the step.
    """

    def __init__(self, layers: Sequence["I64Block"], max_tokens: int, chunks=4, out_dtype=torch.bfloat16):
        if not layers:
            raise ValueError("no layers")
        self.layers = list(layers)
        self.device = self.layers[0].device
        self.hidden = self.layers[0].cfg.hidden_size
        self.weights = [1.0] * max(1, int(chunks)) if isinstance(chunks, int) else [float(w) for w in chunks]
        if not self.weights or min(self.weights) <= 0:
            raise ValueError("micro-batch sizes must be positive")
        self.chunks = len(self.weights)
        self.max_tokens = int(max_tokens)
        # largest micro-batch of a full step, plus one rounding quantum: a shorter step's inner cuts land on other multiples of 256
        self.per = (max(hi - lo for lo, hi in self._cuts(self.max_tokens)) + 256) if self.max_tokens else 1
        dev, h = self.device, self.hidden
        self.x = [torch.empty(self.per, h, dtype=torch.bfloat16, device=dev) for _ in range(2)]
        self.ids = [torch.empty(self.per, dtype=torch.int64, device=dev) for _ in range(2)]
        self.out = [[torch.empty(self.per, h, dtype=out_dtype, device=dev) for _ in range(2)] for _ in range(2)]
        self.h2d = torch.cuda.Stream(dev)
        self.d2h = torch.cuda.Stream(dev)
        self._computed, self._copied, self._seq = [None, None], [None, None], 0     # slot events (carried across run_stream calls)

    def _cuts(self, n: int):
        """[lo, hi) of every micro-batch of an n-token step: proportional to the weights, inner cuts on multiples of 256."""
        tot, acc, cuts, lo = sum(self.weights), 0.0, [], 0
        for i, w in enumerate(self.weights):
            acc += w
            hi = n if i == self.chunks - 1 else min(n, int(round(n * acc / tot / 256.0)) * 256)
            hi = max(hi, lo)
            cuts.append((lo, hi))
            lo = hi
        return cuts

    def run(self, x_host: torch.Tensor, ids_host: Optional[torch.Tensor], out_host: torch.Tensor,
            chain: bool = False) -> None:
        """One self-contained step: the copy streams first wait for the caller's stream, and the caller's stream waits for
        the last copy-out before the call returns to it (capturable in a CUDA graph)."""
        self._step(x_host, ids_host, out_host, chain, streaming=False)

    def run_stream(self, x_host: torch.Tensor, ids_host: Optional[torch.Tensor], out_host: torch.Tensor,
                   chain: bool = False) -> None:
        """A step of a STREAM of steps: nothing joins at the step boundary, so the copy-in of the next step's first micro-batch
        runs under this step's last kernels and this step's last copy-out under the next step's first kernels — the slots'
        events carry over from call to call.  The host buffers must be ready when the call is made (they are read by the copy
        engine, not ordered behind work on the caller's stream) and ``out_host`` is complete only after :meth:`drain`.  Not for
        CUDA-graph capture (use :meth:`run`)."""
        self._step(x_host, ids_host, out_host, chain, streaming=True)

    def drain(self) -> None:
        """The caller's stream waits for every copy issued so far."""
        with torch.cuda.device(self.device):
            cur = torch.cuda.current_stream(self.device)
            cur.wait_stream(self.h2d)
            cur.wait_stream(self.d2h)

    def _step(self, x_host, ids_host, out_host, chain, streaming) -> None:
        n = x_host.shape[0]
        if n > self.max_tokens:
            raise ValueError(f"{n} tokens exceed the pipeline's capacity of {self.max_tokens}")
        if x_host.dtype != torch.bfloat16 or out_host.dtype != self.out[0][0].dtype:
            raise TypeError("x_host must be bfloat16 and out_host of the pipeline's output dtype")
        cuts = self._cuts(n)
        if n and max(hi - lo for lo, hi in cuts) > self.per:
            raise ValueError("micro-batch larger than the pipeline's buffers")
        with torch.cuda.device(self.device):
            cur = torch.cuda.current_stream(self.device)
            if not streaming:
                # steps streamed before this one and never drained: their last copy-outs may still be reading the slots'
                # output buffers (and their kernels the input slots) — order this step behind them before forgetting the events
                for ev in (*self._copied, *self._computed):
                    if ev is not None:
                        cur.wait_event(ev)
                self.h2d.wait_stream(cur)
                self.d2h.wait_stream(cur)
                self._computed, self._copied, self._seq = [None, None], [None, None], 0
            for lo, hi in cuts:
                m = hi - lo
                if m == 0:
                    continue
                s = self._seq & 1
                self._seq += 1
                with torch.cuda.stream(self.h2d):
                    if self._computed[s] is not None:
                        self.h2d.wait_event(self._computed[s])        # the kernels that last read this slot are done
                    self.x[s][:m].copy_(x_host[lo:hi], non_blocking=True)
                    if ids_host is not None:
                        self.ids[s][:m].copy_(ids_host[lo:hi], non_blocking=True)
                    arrived = torch.cuda.Event()
                    arrived.record(self.h2d)
                cur.wait_event(arrived)
                if self._copied[s] is not None:
                    cur.wait_event(self._copied[s])                   # this slot's output buffers have left for the host
                src = self.x[s][:m]
                ids = self.ids[s][:m] if ids_host is not None else None
                last = None
                for i, blk in enumerate(self.layers):
                    last = self.out[s][i & 1][:m]
                    blk.forward(src, ids, last.dtype, out=last)
                    if chain:
                        src = last
                ev = torch.cuda.Event()
                ev.record(cur)
                self._computed[s] = ev
                with torch.cuda.stream(self.d2h):
                    self.d2h.wait_event(ev)
                    out_host[lo:hi].copy_(last, non_blocking=True)
                    ev2 = torch.cuda.Event()
                    ev2.record(self.d2h)
                    self._copied[s] = ev2
            if not streaming:
                cur.wait_stream(self.h2d)
                cur.wait_stream(self.d2h)
                self._computed, self._copied, self._seq = [None, None], [None, None], 0
