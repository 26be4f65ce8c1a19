"""Expert-parallel exchange logic (vllm_i64.parallel.expert_parallel) on 2 and 4 CPU ranks over gloo, with the
expert compute injected from the oracle: the per-rank outputs must equal the single-rank block on the same tokens
and the virtual-rank restatement of the reference's ``_forward_ep`` (layers/token_routed_i64/layer.py:209-298)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import i64_test_helpers as Hh
from oracle import i64_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vllm_i64.parallel.expert_parallel import ep_moe_forward, init_expert_parallel, local_expert_range
        ep = init_expert_parallel()
        h, e, ie, s, v, k, n_total = case
        cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n_total, seed=3)
        cuts = np.linspace(0, n_total, world + 1).astype(int)
        if rank == world - 1 and world > 1:
            pass
        lo, hi = cuts[rank], cuts[rank + 1]
        xl, il = x[lo:hi], ids[lo:hi]
        routes = O.route_table(il, table, hi - lo, e)
        first, local_e = local_expert_range(e, ep)
        gate_up = np.concatenate([w.gate, w.up], axis=2)

        def local_fn(rows, local_ids):
            y = O.fused_token_routed_fp32(rows.numpy(), gate_up[first:first + local_e], w.down[first:first + local_e],
                                          local_ids.numpy().astype(np.int64), local_e)
            return torch.from_numpy(y)

        def shared_fn(xx):
            return torch.from_numpy(O.shared_forward_fp32(xx.numpy(), w))

        def combine(y, shared):
            yk = y.numpy().reshape(k, hi - lo, h)
            return torch.from_numpy(O.combine_fp32([yk[r] for r in range(k)], None if shared is None else shared.numpy(),
                                                   cfg))

        out = ep_moe_forward(torch.from_numpy(xl), torch.from_numpy(routes), e, local_fn, combine,
                             shared_fn if s else None, ep)
        ref = O.moe_block_fp32(xl, il, table, w, cfg)
        ret[rank] = float(np.abs(out.numpy() - ref).max() / max(np.abs(ref).max(), 1e-30))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,case", [(2, (64, 4, 16, 32, 300, 2, 101)), (4, (64, 8, 16, 0, 300, 1, 64)),
                                        (2, (64, 4, 16, 32, 300, 2, 3))])
def test_ep_forward_matches_single_rank(world, case):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), case, ret), nprocs=world, join=True)
    assert len(ret) == world
    for r in range(world):
        assert ret[r] < 1e-5, (r, ret[r])


def test_exchange_plan_is_stable_by_rank():
    """Single-process check of the send order (stable by destination rank, then assignment index) against the
    virtual-rank oracle's ordering."""
    rng = np.random.RandomState(0)
    routes = rng.randint(0, 8, size=(2, 50))
    dest = routes.reshape(-1) // 2
    order = np.argsort(dest, kind="stable")
    t = torch.argsort(torch.from_numpy(dest), stable=True).numpy()
    np.testing.assert_array_equal(order, t)
    assert (np.diff(dest[order]) >= 0).all()


def test_peer_layout_unequal_batches_virtual_ranks():
    """Index arithmetic of the peer-memory exchange (csrc/i64b_ep.cu) restated in numpy over virtual ranks: every source
    plans only its own tokens, its slab for owner d goes into region `src` of d's receive buffer, the owner's group table
    is one group per (local expert, source) built from the published counts.  Ranks bring different numbers of tokens.
    Running the experts group by group on the receive buffers and pulling (owner, row) back must equal the one-rank block."""
    h, e, ie, v, k = 32, 8, 16, 500, 2
    world = 4
    local_e = e // world
    n_per = [37, 0, 64, 5]
    cfg, w, table, ids, x = Hh.make_block(h, e, ie, 0, v, k, sum(n_per), seed=5)
    cap_src = max(n_per) * k
    recv = [np.zeros((world * cap_src, h), np.float32) for _ in range(world)]
    counts = np.zeros((world, world, local_e), np.int64)          # [owner][src][local expert]
    locs, cuts = [], np.concatenate([[0], np.cumsum(n_per)])
    for src in range(world):
        lo, hi = cuts[src], cuts[src + 1]
        routes = O.route_table(ids[lo:hi], table, hi - lo, e)      # [K, n]
        plan, _ = O.merged_layout(routes, e)                       # merged expert-major layout of THIS rank's tokens
        gs = np.concatenate([[0], np.cumsum(plan.group_rows)])
        loc = np.zeros((k, hi - lo, 2), np.int64)
        for kk in range(k):
            for t in range(hi - lo):
                ex = routes[kk, t]
                d = ex // local_e
                row = src * cap_src + plan.dest_row[kk, t] - gs[d * local_e]
                recv[d][row] = x[lo + t]
                loc[kk, t] = (d, row)
        for ex in range(e):
            counts[ex // local_e, src, ex % local_e] = plan.group_rows[ex]
        locs.append(loc)
    ys = []
    for d in range(world):
        y = np.zeros_like(recv[d])
        for le in range(local_e):
            for src in range(world):
                start = src * cap_src + counts[d, src, :le].sum()
                rows = counts[d, src, le]
                ex = d * local_e + le
                seg = recv[d][start:start + rows]
                y[start:start + rows] = O._mm(O._silu(O._mm(seg, w.gate[ex])) * O._mm(seg, w.up[ex]), w.down[ex])
        ys.append(y)
    ref = O.moe_block_fp32(x, ids, table, w, cfg)
    wts = cfg.route_weights()
    for src in range(world):
        lo, hi = cuts[src], cuts[src + 1]
        out = np.zeros((hi - lo, h), np.float32)
        for kk in range(k):
            for t in range(hi - lo):
                d, row = locs[src][kk, t]
                out[t] += np.float32(wts[kk]) * ys[d][row]
        out *= np.float32(cfg.routed_output_scale)
        np.testing.assert_allclose(out, ref[lo:hi], rtol=1e-5, atol=1e-6)
