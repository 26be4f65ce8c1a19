"""Expert parallelism for the token-routed MoE block (BASELINE config 5).

Behavioural spec: ``TokenRoutedMLP._forward_ep`` of the reference's vLLM-fork copy
(layers/token_routed_i64/layer.py:209-298 — not importable in the reference tree): experts are block-partitioned
over the EP group (``rank = expert_id // local_E``, :227), tokens stay data-parallel, one all-to-all carries the
dispatched rows to the owning ranks and one carries the expert outputs back.  Differences from the reference:

  * top-k routes (the reference EP path is top-1 only) — every (route, token) assignment is exchanged;
  * the deterministic dispatch plan (stable by destination rank, then assignment index) replaces argsort;
  * expert ids ride along as int32; the shared expert and the weighted combine run on the token's home rank, so the
    shared expert's GEMMs overlap the dispatch exchange on a side stream;
  * local expert compute is ONE grouped-GEMM block call (``I64Block.forward_routed``), not a per-expert loop.

The exchange uses ``torch.distributed.all_to_all_single`` (NCCL over NVLink/NVSwitch on GPUs; gloo in the CPU unit
tests, where the expert compute is injected).  Split sizes are exchanged first, like the reference (:230-247).
"""
from __future__ import annotations

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


def check_equal_local_batches(n: int, ep: Optional[EPState] = None, device=None) -> None:
    """The peer-memory exchange indexes token t of rank r as ``r*n + t`` in the global plan and all-gathers the ids with
    equal contributions, so unequal local batches would hang in the collective.  Collective call (one small
    all-reduce): raises the same ``ValueError`` on every rank when the ranks disagree on ``n``.  The engine pads decode
    batches to a common size; the reference's own EP path exchanges split sizes through the host instead
    (layers/token_routed_i64/layer.py:230-247)."""
    ep = ep or _EP
    if ep.ep_size <= 1:
        return
    nn = torch.tensor([n, -n], dtype=torch.int64, device=device)
    dist.all_reduce(nn, op=dist.ReduceOp.MAX, group=ep.group)
    n_max, n_min = int(nn[0]), -int(nn[1])
    if n_max != n_min:
        raise ValueError(f"expert-parallel ranks must bring the same number of tokens per step "
                         f"(min {n_min}, max {n_max}): pad the local batch to a common size")


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


class EPTokenRoutedBlock:
    """The block with its routed experts sharded over the EP group (GPU).  Holds two C-ABI blocks: the local
    experts (E/G of them, no shared expert) and the replicated dense shared expert."""

    def __init__(self, hidden_size: int, num_experts: int, expert_inter: int, shared_inter: int, vocab_size: int,
                 top_k: int, route_weights: Sequence[float], routed_scale: float, shared_scale: float, mode: int,
                 device: torch.device, ep: Optional[EPState] = None, transport_dtype: torch.dtype = torch.float32):
        """``transport_dtype``: dtype the expert outputs travel back in.  float32 (default) makes the EP result
        bit-identical to the single-GPU block; bfloat16 is the reference's choice (its EP path returns expert outputs
        in the model dtype, layers/token_routed_i64/layer.py:283-292) and halves the return traffic."""
        from vllm_i64.kernels.block import I64Block
        if transport_dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("transport_dtype must be float32 or bfloat16")
        self.transport_dtype = transport_dtype
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
        self._table = None

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
    def _setup_peers(self, n: int) -> None:
        """Allocate this rank's receive / home buffers for ``n`` local tokens, exchange the CUDA IPC handles and map
        every peer's buffers.  Setup only (allocates and synchronises); ``n`` must be the same on all ranks."""
        import ctypes as C
        from vllm_i64 import _C
        lib = _C.lib()
        G, K, H = self.ep.ep_size, self.K, self.H
        if G > _C.EP_MAX_RANKS:
            raise ValueError(f"at most {_C.EP_MAX_RANKS} expert-parallel ranks")
        check_equal_local_batches(n, self.ep, self.device)
        cap = G * n * K                                        # worst case: every assignment lands here (nothing dropped)

        def up(v):
            return (v + 1023) & ~1023
        off_xs, off_y = 0, up(cap * H * 2)
        off_flags = off_y + up(cap * H * 4)
        total = off_flags + 1024
        with torch.cuda.device(self.device):
            torch.cuda.synchronize(self.device)
            self._release_peers()
            base = C.c_void_p()
            _C.check(lib.i64b_ep_alloc(total, C.byref(base)))
            handle = C.create_string_buffer(64)
            _C.check(lib.i64b_ipc_export(base, handle))
            handles = [None] * G
            if G > 1:
                dist.all_gather_object(handles, bytes(handle.raw), group=self.ep.group)
            bases = []
            for d in range(G):
                if d == self.ep.ep_rank:
                    bases.append(base.value)
                else:
                    p = C.c_void_p()
                    _C.check(lib.i64b_ipc_import(handles[d], C.byref(p)))
                    bases.append(p.value)
        ctx = _C.EPCtx()
        ctx.rank, ctx.size, ctx.n_local, ctx.hidden = self.ep.ep_rank, G, n, H
        ctx.top_k, ctx.num_experts, ctx.vocab, ctx.cap_rows = K, self.E, self.router.cfg.vocab_size, cap
        ctx.y_dtype = _C.F32 if self.transport_dtype == torch.float32 else _C.BF16
        for d in range(G):
            ctx.peer_xs[d] = bases[d] + off_xs
            ctx.peer_y[d] = bases[d] + off_y
            ctx.peer_flags[d] = bases[d] + off_flags
        ctx.state = base.value + off_flags + 512
        self._ctx, self._bases, self._n_setup = ctx, bases, n
        self._ws_plan = torch.empty(lib.i64b_ep_workspace_bytes(C.byref(ctx)) + 256, dtype=torch.uint8, device=self.device)
        self._ws_gemm = torch.empty(lib.i64b_moe_dispatched_workspace_bytes(self.local._h, C.byref(ctx)) + 256,
                                    dtype=torch.uint8, device=self.device)
        self._ids_all = torch.empty(G * n, dtype=torch.int64, device=self.device)
        # persistent fp32 output of the replicated shared expert (a fresh 100+ MB torch allocation per call was
        # measured at ~350 us of allocator time per layer)
        self._shared_out = (torch.empty(n, H, dtype=torch.float32, device=self.device)
                            if self.shared is not None else None)
        if G > 1:
            dist.barrier(group=self.ep.group)                  # every rank has mapped every buffer

    def _release_peers(self) -> None:
        from vllm_i64 import _C
        bases = getattr(self, "_bases", None)
        if not bases:
            return
        lib = _C.lib()
        for d, b in enumerate(bases):
            if d == self.ep.ep_rank:
                lib.i64b_ep_free(b)
            else:
                lib.i64b_ipc_close(b)
        self._bases = None

    def forward_peer(self, x: torch.Tensor, token_ids: torch.Tensor, out_dtype=torch.bfloat16,
                     out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One expert-parallel forward over NVLink peer memory: no host synchronisation, no NCCL on the data path
        (NCCL carries only the all-gather of the token ids).  See include/i64b.h (expert parallelism)."""
        import ctypes as C
        from vllm_i64 import _C
        _C.require_cuda(x, token_ids)
        lib = _C.lib()
        n = x.shape[0]
        if x.dtype != torch.bfloat16 or not x.is_contiguous():
            x = x.to(torch.bfloat16).contiguous()
        token_ids = token_ids.to(torch.int64).contiguous()
        if getattr(self, "_n_setup", -1) != n:
            self._setup_peers(n)
        ctx = C.byref(self._ctx)
        with torch.cuda.device(self.device):
            cur = torch.cuda.current_stream(self.device)
            shared = None
            if self.shared is not None:
                # the replicated shared expert runs on a side stream: its GEMMs fill the SMs while the dispatch kernel
                # (NVLink-bound, a few warps per SM) and the barrier wait occupy the main stream
                if getattr(self, "_side", None) is None:
                    self._side = torch.cuda.Stream(self.device)
                self._side.wait_stream(cur)
                with torch.cuda.stream(self._side):
                    shared = self.shared.forward(x, None, torch.float32, out=self._shared_out)
            st = _C.stream_ptr(self.device)
            if self.ep.ep_size > 1:
                dist.all_gather_into_tensor(self._ids_all, token_ids, group=self.ep.group)
            else:
                self._ids_all.copy_(token_ids)
            _C.check(lib.i64b_ep_dispatch(ctx, x.data_ptr(), self._ids_all.data_ptr(), self._table.data_ptr(),
                                          self.local.cfg.expert_inter, self._ws_plan.data_ptr(), self._ws_plan.numel(), st))
            _C.check(lib.i64b_ep_signal(ctx, st))
            _C.check(lib.i64b_ep_wait(ctx, st))
            _C.check(lib.i64b_moe_forward_dispatched(self.local._h, ctx, self._ws_plan.data_ptr(), self._ws_gemm.data_ptr(),
                                                     self._ws_gemm.numel(), st))
            _C.check(lib.i64b_ep_signal(ctx, st))
            _C.check(lib.i64b_ep_wait(ctx, st))
            if out is None:
                out = torch.empty(n, self.H, dtype=out_dtype, device=self.device)
            if shared is not None:
                cur.wait_stream(self._side)
            wts = (C.c_float * 4)(*[self.w[i] if i < len(self.w) else 0.0 for i in range(4)])
            _C.check(lib.i64b_ep_combine(ctx, self._ws_plan.data_ptr(), _C.ptr(shared), wts, self.rs, self.ss,
                                         out.data_ptr(), _C.dtype_code(out), st))
        return out

    def forward(self, x: torch.Tensor, token_ids: torch.Tensor, out_dtype=torch.bfloat16,
                exchange: str = "peer", out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``exchange="peer"``: NVLink peer-memory path (default).  ``"nccl"``: all_to_all_single exchange with
        host-side split sizes, the reference's scheme — kept as the comparison baseline (bench.py reports both)."""
        if exchange == "peer":
            return self.forward_peer(x, token_ids, out_dtype, out)
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
