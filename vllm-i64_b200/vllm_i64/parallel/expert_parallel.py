"""Expert parallelism for the token-routed MoE block (BASELINE config 5).

Behavioural spec: ``TokenRoutedMLP._forward_ep`` of the reference's vLLM-fork copy
(layers/token_routed_i64/layer.py:209-298 — not importable in the reference tree): experts are block-partitioned
over the EP group (``rank = expert_id // local_E``, :227), tokens stay data-parallel, dispatched rows travel to the
owning ranks and the expert outputs travel back.

Product path (``EPTokenRoutedBlock.forward_peer``, csrc/i64b_ep.cu): NVLink peer memory, no collective library call on
the data path.  A rank plans only its own tokens; its dispatch kernel writes every row straight into "its" region of
the owning rank's receive buffer and publishes per-expert row counts in the owner's flags page; a flag barrier in peer
memory replaces the collective's implicit synchronisation; the owners run their experts as one grouped GEMM pair; the
home rank's combine kernel pulls the K rows of each token back over NVLink.  The shared expert and the weighted
combine run on the token's home rank, the shared expert on a second stream.  Ranks may bring different numbers of
tokens; nothing synchronises with the host, so the whole step is CUDA-graph capturable.

Baseline path (``exchange="nccl"``, and ``ep_moe_forward`` for the CPU/gloo unit tests): the reference's scheme —
split sizes exchanged through the host, ``all_to_all_single`` each way — generalised to top-k routes.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist


@dataclass
class EPState:
    ep_size: int = 1
    ep_rank: int = 0
    group: Optional[dist.ProcessGroup] = None


_EP = EPState()


def get_ep() -> EPState:
    return _EP


def init_expert_parallel(ep_size: Optional[int] = None, group: Optional[dist.ProcessGroup] = None) -> EPState:
    """EP group = all ranks of ``group`` (default: the world).  ``dist`` must be initialised."""
    global _EP
    if not dist.is_initialized():
        _EP = EPState()
        return _EP
    size = dist.get_world_size(group)
    if ep_size is not None and ep_size != size:
        raise ValueError(f"ep_size {ep_size} != group size {size}")
    _EP = EPState(ep_size=size, ep_rank=dist.get_rank(group), group=group)
    return _EP


def local_expert_range(num_experts: int, ep: Optional[EPState] = None):
    """Block partition ``e -> rank e // (E / G)`` (layers/token_routed_i64/layer.py:84-88)."""
    ep = ep or _EP
    if num_experts % ep.ep_size != 0:
        raise ValueError(f"num_experts ({num_experts}) must be divisible by ep_size ({ep.ep_size})")
    local = num_experts // ep.ep_size
    return ep.ep_rank * local, local


@dataclass
class ExchangePlan:
    order: torch.Tensor          # [A] assignment index (k*N + t) in send order (stable by destination rank)
    send_counts: List[int]       # rows to each rank
    recv_counts: List[int]       # rows from each rank


def make_exchange_plan(routes: torch.Tensor, num_experts: int, ep: EPState) -> ExchangePlan:
    """routes ``[K,N]`` (expert ids) → deterministic send order + split sizes (counts all-to-all, :230-247)."""
    local = num_experts // ep.ep_size
    dest = (routes.reshape(-1).to(torch.int64) // local)
    order = torch.argsort(dest, stable=True)
    send = torch.bincount(dest, minlength=ep.ep_size).to(torch.int64)
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=ep.group)
    return ExchangePlan(order, send.tolist(), recv.tolist())


def ep_moe_forward(x: torch.Tensor, routes: torch.Tensor, num_experts: int,
                   local_expert_fn: Callable[[torch.Tensor, torch.Tensor], torch.Tensor],
                   combine_fn: Callable[[torch.Tensor, torch.Tensor], torch.Tensor],
                   shared_fn: Optional[Callable[[torch.Tensor], torch.Tensor]] = None,
                   ep: Optional[EPState] = None, return_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """One expert-parallel MoE FFN forward on this rank's tokens.

    ``local_expert_fn(rows [M,H], local_ids [M]) -> [M,H]`` runs this rank's experts on the received rows;
    ``combine_fn(y [K*N,H] in assignment order (k*N+t), shared [N,H] | None) -> [N,H]`` applies the route weights
    and output scales; ``shared_fn(x) -> [N,H]`` is the replicated shared expert (overlaps the exchange).
    """
    ep = ep or _EP
    k, n = routes.shape
    h = x.shape[1]
    plan = make_exchange_plan(routes, num_experts, ep)
    local = num_experts // ep.ep_size
    tok = plan.order % n
    send_x = x.index_select(0, tok)
    send_e = routes.reshape(-1).index_select(0, plan.order).to(torch.int32)
    m = sum(plan.recv_counts)
    recv_x = torch.empty(m, h, dtype=x.dtype, device=x.device)
    recv_e = torch.empty(m, dtype=torch.int32, device=x.device)

    side = None
    shared = None
    if shared_fn is not None and x.is_cuda:
        side = torch.cuda.Stream(x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side):
            shared = shared_fn(x)
    dist.all_to_all_single(recv_x, send_x, plan.recv_counts, plan.send_counts, group=ep.group)
    dist.all_to_all_single(recv_e, send_e, plan.recv_counts, plan.send_counts, group=ep.group)
    if shared_fn is not None and side is None:
        shared = shared_fn(x)

    y_recv = local_expert_fn(recv_x, recv_e - ep.ep_rank * local).to(return_dtype)
    y_back = torch.empty(k * n, h, dtype=return_dtype, device=x.device)
    dist.all_to_all_single(y_back, y_recv.contiguous(), plan.send_counts, plan.recv_counts, group=ep.group)
    y = torch.empty_like(y_back)
    y.index_copy_(0, plan.order, y_back)                    # un-sort to assignment order (layer.py:294-295)
    if side is not None:
        torch.cuda.current_stream(x.device).wait_stream(side)
        shared.record_stream(torch.cuda.current_stream(x.device))
    return combine_fn(y, shared)


class _Lane:
    """One micro-batch pipeline of an EPTokenRoutedBlock: peer-mapped buffers + flags (ctx), workspaces, stream."""
    __slots__ = ("ctx", "bases", "cap", "ws_plan", "ws_gemm", "shared_out", "stream", "dispatched", "consumed")


class EPTokenRoutedBlock:
    """The block with its routed experts sharded over the EP group (GPU).  Holds two C-ABI blocks: the local
    experts (E/G of them, no shared expert) and the replicated dense shared expert."""

    def __init__(self, hidden_size: int, num_experts: int, expert_inter: int, shared_inter: int, vocab_size: int,
                 top_k: int, route_weights: Sequence[float], routed_scale: float, shared_scale: float, mode: int,
                 device: torch.device, ep: Optional[EPState] = None, transport_dtype: torch.dtype = torch.bfloat16):
        """``transport_dtype``: dtype the expert outputs travel back in.  bfloat16 (default) is the reference's choice
        (its EP path returns expert outputs in the model dtype, layers/token_routed_i64/layer.py:283-292) and halves
        the return traffic: one more bf16 rounding per assignment than the single-GPU block.  float32 makes the EP
        result bit-identical to the single-GPU block.  ``mode`` ``_C.MODE_W8A8``: the rows are quantised once on the
        home rank and travel as int8 + one fp32 scale (half the dispatch bytes); the owners run the W8A8 experts."""
        from vllm_i64 import _C
        from vllm_i64.kernels.block import I64Block
        if transport_dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("transport_dtype must be float32 or bfloat16")
        self.transport_dtype = transport_dtype
        self.mode = int(mode)
        self.ep = ep or _EP
        self.E, self.K, self.H = num_experts, top_k, hidden_size
        self.first, self.local_E = local_expert_range(num_experts, self.ep)
        self.w, self.rs, self.ss = list(route_weights), routed_scale, shared_scale
        self.device = torch.device(device)
        self.router = I64Block(hidden_size, num_experts, expert_inter, 0, vocab_size, top_k, route_weights, 1.0, 1.0,
                               mode, device)               # routing table owner (no weights loaded)
        self.local = I64Block(hidden_size, self.local_E, expert_inter, 0, 1, 1, [1.0], 1.0, 1.0, mode, device)
        self.shared = (I64Block(hidden_size, 0, 0, shared_inter, 1, 0, [], 1.0, 1.0, mode, device)
                       if shared_inter > 0 else None)
        if self.shared is not None:
            # the shared expert runs on a second stream next to the exchange kernels: keep it off the single-launch decode
            # kernel, whose CTAs wait on one another and must own the GPU (include/i64b.h, i64b_moe_forward)
            self.shared.set_debug(_C.DEBUG_NO_FUSED)
        self._table = None
        self._lanes = None
        self._bases = None
        self._cap_tokens = 0
        self._planned = None
        self.timeline = None

    def __del__(self):
        # best effort, non-collective: unmap and free without the barrier (the process is going away)
        try:
            from vllm_i64 import _C
            lanes = getattr(self, "_lanes", None)
            if lanes:
                lib = _C.lib()
                for lane in lanes:
                    for d, b in enumerate(lane.bases):
                        (lib.i64b_ep_free if d == self.ep.ep_rank else lib.i64b_ipc_close)(b)
                self._lanes = None
        except Exception:
            pass

    def load_w8a8(self, w13_q, w13_s, w2_q, w2_s, sh13_q=None, sh13_s=None, sh2_q=None, sh2_s=None) -> None:
        """FULL materialised int8 weights (include/i64b.h, i64b_moe_load_w8a8); this rank keeps its experts' slice."""
        lo, hi = self.first, self.first + self.local_E
        self.local.load_w8a8(w13_q[lo:hi], w13_s[lo:hi], w2_q[lo:hi], w2_s[lo:hi])
        if self.shared is not None:
            self.shared.load_w8a8(None, None, None, None, sh13_q, sh13_s, sh2_q, sh2_s)

    def load_bf16(self, gate, up, down, shared_gate=None, shared_up=None, shared_down=None) -> None:
        """FULL reference-layout weights; this rank keeps experts [first, first + local_E) and the shared expert."""
        lo, hi = self.first, self.first + self.local_E
        self.local.load_bf16(gate[lo:hi], up[lo:hi], down[lo:hi])
        if self.shared is not None:
            self.shared.load_bf16(None, None, None, shared_gate, shared_up, shared_down)

    def set_table(self, table_i64: torch.Tensor) -> None:
        self.router.set_table(table_i64)
        self._table = self.router._table_u8

    def route(self, token_ids: torch.Tensor) -> torch.Tensor:
        from vllm_i64 import _C
        n = token_ids.numel()
        out = torch.empty(self.K, n, dtype=torch.int64, device=self.device)
        _C.check(_C.lib().i64b_route_table(token_ids.data_ptr(), self._table.data_ptr(), out.data_ptr(), n,
                                           self.router.cfg.vocab_size, self.K, self.E, _C.stream_ptr(self.device)))
        return out

    # ------------------------------------------------------------------ NVLink peer-memory path (the product path)
    def reserve(self, max_tokens: int, lanes: Optional[int] = None) -> None:
        """COLLECTIVE, setup only (allocates and synchronises): size this rank's receive / output buffers for
        ``max_tokens`` local tokens per step on EVERY rank (the maximum over the ranks is used), exchange the CUDA IPC
        handles once and map every peer's buffers.  The engine calls it with its maximum batch; ``forward_peer`` calls
        it on first use.

        ``lanes``: a forward cuts its tokens into that many micro-batches, each with its own buffers, flags and stream,
        started one dispatch apart — while one lane's rows cross NVLink the other lane's GEMMs own the tensor cores.
        Every rank must use the same number.  Default 1 (env ``I64B_EP_LANES`` overrides): measured on B200 at 2 and 8
        GPUs the pipeline LOSES (config 5, 64K tokens: 193 us/layer with one lane, 271-299 us with two) — the exchange
        kernels are fast only with all SMs behind them (a 16-SM partition pulls at 230 GB/s, all SMs at 520), and twice
        as many half-sized GEMMs and barriers cost more than the overlap returns.  Kept because it is exact and tested."""
        import ctypes as C
        from vllm_i64 import _C
        lib = _C.lib()
        G, K, H = self.ep.ep_size, self.K, self.H
        if G > _C.EP_MAX_RANKS:
            raise ValueError(f"at most {_C.EP_MAX_RANKS} expert-parallel ranks")
        n = int(max_tokens)
        if G > 1:
            nn = torch.tensor([n], dtype=torch.int64, device=self.device)
            dist.all_reduce(nn, op=dist.ReduceOp.MAX, group=self.ep.group)
            n = int(nn[0])
        n = max(n, 1)
        if lanes is None:
            lanes = int(os.environ.get("I64B_EP_LANES", "0")) or 1
        lanes = max(1, min(int(lanes), 8))
        n_lane = -(-n // lanes)
        # SM partition: the exchange kernels get `comm_sms` SMs of their own while GEMMs run next to them (the shared
        # expert next to the dispatch; with several lanes, every GEMM next to the other lanes' exchange).  A GEMM CTA
        # takes ~210 KB of shared memory: without the partition the exchange kernels would queue behind whole GEMMs.
        sms = torch.cuda.get_device_properties(self.device).multi_processor_count
        comm = int(os.environ.get("I64B_EP_COMM_SMS", "16")) if G > 1 else 0
        comm = max(0, min(comm, sms // 2))
        self._comm_dispatch = comm if (self.shared is not None or lanes > 1) else 0
        self._comm_combine = comm if lanes > 1 else 0
        if self.shared is not None:
            self.shared.set_cta_limit(sms - comm if self._comm_dispatch else 0)
        self.local.set_cta_limit(sms - comm if (comm and lanes > 1) else 0)
        cap_src = n_lane * K                                   # a source's whole micro-batch may land on one owner
        rows = G * cap_src
        q = self.mode == _C.MODE_W8A8
        ysz = 4 if self.transport_dtype == torch.float32 else 2

        def up(v):
            return (v + 1023) & ~1023
        off_xs = 0
        off_scale = up(rows * H * (1 if q else 2))
        off_y = off_scale + up(rows * 4)
        off_flags = off_y + up(rows * H * ysz)
        off_state = off_flags + up(_C.EP_FLAGS_INTS * 4)
        total = off_state + 1024
        with torch.cuda.device(self.device):
            torch.cuda.synchronize(self.device)
            self.close()
            mine = []
            for _ in range(lanes):
                base = C.c_void_p()
                _C.check(lib.i64b_ep_alloc(total, C.byref(base)))
                handle = C.create_string_buffer(64)
                _C.check(lib.i64b_ipc_export(base, handle))
                mine.append((base, bytes(handle.raw)))
            handles = [None] * G
            if G > 1:
                dist.all_gather_object(handles, [h for _, h in mine], group=self.ep.group)
            self._lanes = []
            for li in range(lanes):
                base = mine[li][0]
                bases = []
                for d in range(G):
                    if d == self.ep.ep_rank:
                        bases.append(base.value)
                    else:
                        p = C.c_void_p()
                        _C.check(lib.i64b_ipc_import(handles[d][li], C.byref(p)))
                        bases.append(p.value)
                ctx = _C.EPCtx()
                ctx.rank, ctx.size, ctx.n_local, ctx.hidden = self.ep.ep_rank, G, 0, H
                ctx.top_k, ctx.num_experts, ctx.vocab, ctx.cap_src = K, self.E, self.router.cfg.vocab_size, cap_src
                ctx.y_dtype = _C.F32 if self.transport_dtype == torch.float32 else _C.BF16
                ctx.x_mode = self.mode
                for d in range(G):
                    ctx.peer_xs[d] = bases[d] + off_xs
                    ctx.peer_xs_scale[d] = bases[d] + off_scale
                    ctx.peer_y[d] = bases[d] + off_y
                    ctx.peer_flags[d] = bases[d] + off_flags
                ctx.state = base.value + off_state
                lane = _Lane()
                lane.ctx, lane.bases, lane.cap = ctx, bases, n_lane
                lane.ws_plan = torch.empty(lib.i64b_ep_workspace_bytes(C.byref(ctx)) + 256, dtype=torch.uint8,
                                           device=self.device)
                lane.ws_gemm = torch.empty(lib.i64b_moe_dispatched_workspace_bytes(self.local._h, C.byref(ctx)) + 256,
                                           dtype=torch.uint8, device=self.device)
                # persistent fp32 output of the home rank's shared expert (a fresh 100+ MB torch allocation per call
                # was measured at ~350 us of allocator time per layer)
                lane.shared_out = (torch.empty(n_lane, H, dtype=torch.float32, device=self.device)
                                   if self.shared is not None else None)
                lane.stream = torch.cuda.Stream(self.device)
                lane.dispatched = torch.cuda.Event()
                lane.consumed = None            # recorded after each forward's combine: the last reader of the lane's plan
                self._lanes.append(lane)
        self._cap_tokens = n_lane * lanes
        l0 = self._lanes[0]
        # single-lane names (tools/ep_probe.py, bring-up)
        self._ctx, self._bases, self._ws_plan, self._ws_gemm = l0.ctx, l0.bases, l0.ws_plan, l0.ws_gemm
        self._shared_out, self._side = l0.shared_out, l0.stream
        if G > 1:
            dist.barrier(group=self.ep.group)                  # every rank has mapped every buffer

    def close(self) -> None:
        """COLLECTIVE when buffers are mapped: unmap the peers' buffers, wait until every rank has done so, free ours."""
        from vllm_i64 import _C
        lanes = getattr(self, "_lanes", None)
        if not lanes:
            return
        lib = _C.lib()
        torch.cuda.synchronize(self.device)
        for lane in lanes:
            for d, b in enumerate(lane.bases):
                if d != self.ep.ep_rank:
                    lib.i64b_ipc_close(b)
        if self.ep.ep_size > 1 and dist.is_initialized():
            dist.barrier(group=self.ep.group)                  # nobody still maps the buffers that are freed next
        for lane in lanes:
            lib.i64b_ep_free(lane.bases[self.ep.ep_rank])
        self._lanes = None
        self._bases = None
        self._cap_tokens = 0

    def plan_ahead(self, token_ids: torch.Tensor, stream: Optional[torch.cuda.Stream] = None,
                   step: Optional[int] = None) -> None:
        """Compute this layer's dispatch plan for ``token_ids`` now, on ``stream`` (default: the current one), ahead of
        the ``forward`` that will use it.  Routing depends on the token ids only, so an engine can plan every layer of
        a step on a side stream while earlier layers run (include/i64b.h, i64b_ep_plan); the next ``forward`` with the
        SAME ``token_ids`` tensor waits for the plan and only moves rows.  Single-lane path only (a no-op otherwise).

        Ordering: the plan overwrites the lane's plan workspace, which this layer's PREVIOUS forward reads up to its
        combine; ``stream`` is therefore made to wait for that combine here (no caller-side wait needed for that).  The
        ids themselves must be complete on ``stream`` — if another stream writes them, the caller orders that.

        Validity: the plan is used by the next forward only if that forward brings the same memory (this call keeps a
        reference to ``token_ids``, so the allocator cannot hand the address to another tensor meanwhile), the same
        count, an unchanged in-place version (``copy_`` / ``fill_`` into a static ids buffer after this call discard the
        plan) and the same ``step``.  ``step`` is the caller's generation token (engine step number): pass it to both
        calls when the ids buffer is written by something PyTorch's version counter does not see (a raw kernel, a graph
        replay, ``.data``).  Any mismatch drops the plan and the forward plans for itself."""
        import ctypes as C
        from vllm_i64 import _C
        n = token_ids.numel()
        if not getattr(self, "_lanes", None):
            self.reserve(n)
        if len(self._lanes) != 1 or n == 0:
            return
        if n > self._cap_tokens:
            raise ValueError(f"{n} tokens exceed the reserved expert-parallel capacity of {self._cap_tokens}")
        if token_ids.dtype != torch.int64 or not token_ids.is_contiguous():
            raise TypeError("plan_ahead: contiguous int64 token ids")
        lane = self._lanes[0]
        lane.ctx.n_local = n
        with torch.cuda.device(self.device):
            s = stream or torch.cuda.current_stream(self.device)
            if lane.consumed is not None:
                s.wait_event(lane.consumed)         # the previous forward of this layer has finished reading its plan
            _C.check(_C.lib().i64b_ep_plan(C.byref(lane.ctx), token_ids.data_ptr(), self._table.data_ptr(),
                                           lane.ws_plan.data_ptr(), lane.ws_plan.numel(), s.cuda_stream))
            ev = torch.cuda.Event()
            ev.record(s)
        self._planned = (token_ids, n, token_ids._version, step, ev)

    def forward_peer(self, x: torch.Tensor, token_ids: torch.Tensor, out_dtype=torch.bfloat16,
                     out: Optional[torch.Tensor] = None, step: Optional[int] = None) -> torch.Tensor:
        """One expert-parallel forward over NVLink peer memory: no host synchronisation and no collective call
        (graph-capturable once ``reserve`` has run).  See include/i64b.h (expert parallelism)."""
        import ctypes as C
        from vllm_i64 import _C
        _C.require_cuda(x, token_ids)
        lib = _C.lib()
        n = x.shape[0]
        if x.dtype != torch.bfloat16 or not x.is_contiguous():
            x = x.to(torch.bfloat16).contiguous()
        token_ids = token_ids.to(torch.int64).contiguous()
        if not getattr(self, "_lanes", None):
            self.reserve(n)
        if n > self._cap_tokens:
            raise ValueError(f"{n} tokens exceed the reserved expert-parallel capacity of {self._cap_tokens} tokens per "
                             f"rank: call reserve(max_tokens) on every rank first")
        if out is None:
            out = torch.empty(n, self.H, dtype=out_dtype, device=self.device)
        wts = (C.c_float * 4)(*[self.w[i] if i < len(self.w) else 0.0 for i in range(4)])
        lanes = self._lanes
        with torch.cuda.device(self.device):
            cur = torch.cuda.current_stream(self.device)
            if len(lanes) > 1:
                return self._forward_lanes(x, token_ids, out, wts, cur)
            lane = lanes[0]
            lane.ctx.n_local = n
            ctx = C.byref(lane.ctx)
            shared = None
            if self.shared is not None and n > 0:
                # the home rank's shared expert runs on a side stream: its GEMMs fill the SMs while the dispatch kernel
                # (NVLink-bound, a few warps per SM) and the barrier occupy the main stream
                lane.stream.wait_stream(cur)
                with torch.cuda.stream(lane.stream):
                    shared = self.shared.forward(x, None, torch.float32, out=lane.shared_out[:n])
            st = _C.stream_ptr(self.device)
            tl = self.timeline          # bring-up: a list collects (stage, event) pairs of this call (eager runs only)

            def mark(name):
                if tl is not None:
                    ev = torch.cuda.Event(enable_timing=True)
                    ev.record(cur)
                    tl.append((name, ev))
            mark("start")
            lane.ctx.comm_ctas = self._comm_dispatch
            ids_ptr = token_ids.data_ptr()
            planned, self._planned = self._planned, None
            if (planned is not None and n > 0 and planned[0].data_ptr() == ids_ptr and planned[1] == n
                    and planned[2] == token_ids._version and planned[3] == step):
                cur.wait_event(planned[4])                 # plan_ahead made the plan: rows only
                ids_ptr = None
            elif planned is not None:
                cur.wait_event(planned[4])                 # stale plan: let it finish writing the workspace, then plan afresh
            _C.check(lib.i64b_ep_dispatch(ctx, x.data_ptr(), ids_ptr, self._table.data_ptr(),
                                          lane.ws_plan.data_ptr(), lane.ws_plan.numel(), st))
            mark("plan+dispatch")
            _C.check(lib.i64b_ep_barrier(ctx, st))
            mark("barrier1")
            _C.check(lib.i64b_moe_forward_dispatched(self.local._h, ctx, lane.ws_plan.data_ptr(), lane.ws_gemm.data_ptr(),
                                                     lane.ws_gemm.numel(), st))
            mark("groups+gemm1+gemm2")
            _C.check(lib.i64b_ep_barrier(ctx, st))
            mark("barrier2")
            if shared is not None:
                cur.wait_stream(lane.stream)
            mark("join shared-expert stream")
            lane.ctx.comm_ctas = self._comm_combine
            _C.check(lib.i64b_ep_combine(ctx, lane.ws_plan.data_ptr(), _C.ptr(shared), wts, self.rs, self.ss,
                                         out.data_ptr(), _C.dtype_code(out), st))
            mark("combine (pull)")
            if not torch.cuda.is_current_stream_capturing():
                if lane.consumed is None:
                    lane.consumed = torch.cuda.Event()
                lane.consumed.record(cur)                  # plan_ahead for this layer's next step waits on this
        return out

    def _forward_lanes(self, x, token_ids, out, wts, cur) -> torch.Tensor:
        """Micro-batch pipeline: lane i handles tokens [cuts[i], cuts[i+1]) on its own stream and buffers, in the order
        dispatch -> shared expert (local, covers the wait for the peers' rows) -> barrier -> routed GEMMs -> barrier ->
        combine; lane i+1 starts when lane i's dispatch has been issued, so exchange kernels of one lane run next to GEMMs
        of the other.  Per-token arithmetic does not depend on the cut, so the result equals the single-lane one bit for
        bit.  Every rank runs every lane (an empty micro-batch still publishes its zero counts and joins the barriers)."""
        import ctypes as C
        from vllm_i64 import _C
        lib = _C.lib()
        n = x.shape[0]
        L = len(self._lanes)
        per = -(-n // L) if n else 0
        prev = None
        for li, lane in enumerate(self._lanes):
            lo, hi = min(n, li * per), min(n, (li + 1) * per)
            m = hi - lo
            lane.ctx.n_local = m
            lane.ctx.comm_ctas = self._comm_dispatch       # == _comm_combine with several lanes
            ctx = C.byref(lane.ctx)
            s = lane.stream
            s.wait_stream(cur)
            if prev is not None:
                s.wait_event(prev.dispatched)
            with torch.cuda.stream(s):
                st = s.cuda_stream
                xs, ids = x[lo:hi], token_ids[lo:hi]
                _C.check(lib.i64b_ep_dispatch(ctx, xs.data_ptr() if m else None, ids.data_ptr() if m else None,
                                              self._table.data_ptr(), lane.ws_plan.data_ptr(), lane.ws_plan.numel(), st))
                lane.dispatched.record(s)
                shared = None
                if self.shared is not None and m > 0:
                    shared = self.shared.forward(xs, None, torch.float32, out=lane.shared_out[:m])
                _C.check(lib.i64b_ep_barrier(ctx, st))
                _C.check(lib.i64b_moe_forward_dispatched(self.local._h, ctx, lane.ws_plan.data_ptr(),
                                                         lane.ws_gemm.data_ptr(), lane.ws_gemm.numel(), st))
                _C.check(lib.i64b_ep_barrier(ctx, st))
                o = out[lo:hi]
                _C.check(lib.i64b_ep_combine(ctx, lane.ws_plan.data_ptr(), _C.ptr(shared), wts, self.rs, self.ss,
                                             o.data_ptr() if m else None, _C.dtype_code(out), st))
            prev = lane
        for lane in self._lanes:
            cur.wait_stream(lane.stream)
        return out

    def forward(self, x: torch.Tensor, token_ids: torch.Tensor, out_dtype=torch.bfloat16,
                exchange: str = "peer", out: Optional[torch.Tensor] = None, step: Optional[int] = None) -> torch.Tensor:
        """``exchange="peer"``: NVLink peer-memory path (default).  ``"nccl"``: all_to_all_single exchange with
        host-side split sizes, the reference's scheme — kept as the comparison baseline (bench.py reports both)."""
        if exchange == "peer":
            return self.forward_peer(x, token_ids, out_dtype, out, step)
        from vllm_i64 import _C
        n = x.shape[0]
        routes = self.route(token_ids)

        def local_fn(rows, ids):
            if rows.shape[0] == 0:
                return torch.empty(0, self.H, dtype=torch.float32, device=self.device)
            return self.local.forward_routed(rows, ids.view(1, -1), [1.0], 1.0, 1.0, False, torch.float32)

        def shared_fn(xx):
            return self.shared.forward(xx, None, torch.float32)

        def combine(y, shared):
            out = torch.empty(n, self.H, dtype=out_dtype, device=self.device)
            dest = torch.arange(self.K * n, dtype=torch.int32, device=self.device)      # y is in assignment order
            wts = (_C.C.c_float * 4)(*[self.w[i] if i < len(self.w) else 0.0 for i in range(4)])
            _C.check(_C.lib().i64b_combine(y.data_ptr(), dest.data_ptr(), _C.ptr(shared), self.K, wts, self.rs, self.ss,
                                           out.data_ptr(), _C.dtype_code(out), n, self.H, _C.stream_ptr(self.device)))
            return out

        return ep_moe_forward(x, routes, self.E, local_fn, combine, shared_fn if self.shared else None, self.ep)
