"""Batch-sharded dense InfoNCE with cross-GPU negatives (no reference counterpart).

Each rank owns its images' pixels ("rows").  One all-gather of the L2-normalised bf16 embeddings
(+ class bits) makes every pixel of every rank a key ("cols"); the row-sum pass and the gradient
pass then run on  local rows x all cols.  Because the similarity matrix is symmetric, the
gradient a pixel receives as a *key* of remote queries is sum_j p_ji e_j, which the owning rank can
form itself once it knows every pixel's 1/rowsum — so the reduce-scatter of key gradients
([N_g, D] fp32) that a query-side formulation needs is replaced by an all-gather of N_g floats.
Small all-reduces carry the class sums, the prototype gradient and the loss.

Result contract: loss and d(loss)/d(x_local) at G ranks == the single-GPU op on the concatenated
batch == the CPU oracle on the concatenated batch.  The loss is already the GLOBAL mean: do not
average the returned gradient across ranks again.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from . import _lib
from .ops import _as_feature_operand
from .stages import CudaStages


_SIZES_CHECKED = set()   # (rows_pad, world, group id) already validated across ranks


def _world(group):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def _all_gather_into(out, inp, group):
    try:
        dist.all_gather_into_tensor(out, inp, group=group)
    except (RuntimeError, NotImplementedError):   # backends without the flat form
        g = dist.get_world_size(group)
        dist.all_gather(list(out.chunk(g, dim=0)), inp, group=group)


def _all_reduce_many(tensors, group):
    """Sum all-reduces of several tensors (different dtypes) as ONE NCCL group call."""
    if dist.get_backend(group) == "nccl":
        with dist._coalescing_manager(group=group, device=tensors[0].device):
            for t in tensors:
                dist.all_reduce(t, group=group)
        return
    for t in tensors:
        dist.all_reduce(t, group=group)


def _all_gather_many(pairs, group):
    """All-gathers of several (out, inp) pairs of different dtypes.  On NCCL they are issued as ONE
    group call (ncclGroupStart/End through torch's coalescing manager), so the class bits and tile
    words ride in the launch of the embedding gather instead of paying a collective latency each."""
    if dist.get_backend(group) == "nccl":
        with dist._coalescing_manager(group=group, device=pairs[0][0].device):
            for out, inp in pairs:
                dist.all_gather_into_tensor(out, inp, group=group)
        return
    for out, inp in pairs:
        _all_gather_into(out, inp, group)


def _use_sym_ring(stages, dim):
    if os.environ.get("MUSCLE_ROWSUM_SYM", "1") == "0":
        return False
    return hasattr(stages, "rowsum_sym_ring") and dim in stages.sym_dims()


def sharded_fwd(stages, x, cls, temperature, num_classes, eps, need_grad, group=None):
    """Runs the forward stages on this rank.  -> (loss 0-dim tensor, state for `sharded_bwd` or None).  `stages` is a
    CudaStages (or the CPU stand-in the gloo tests use)."""
    G, rank = _world(group)
    xc, layout, batch, ppi, dim = _as_feature_operand(x)
    n = batch * ppi
    cls_c = cls.reshape(-1).contiguous()
    if cls_c.dtype not in (torch.int64, torch.int32):
        cls_c = cls_c.to(torch.int64)
    if cls_c.numel() != n:
        raise ValueError("cls has %d entries for %d pixels" % (cls_c.numel(), n))
    rows_pad = stages.pad_rows(n)
    if G > 1 and (rows_pad, G, id(group)) not in _SIZES_CHECKED:
        # once per shape and group: the check reads two reduced scalars back (a host sync every step
        # would keep the host from running ahead of the device)
        sizes = torch.tensor([rows_pad], dtype=torch.int64, device=xc.device)
        lo, hi = sizes.clone(), sizes.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        if int(lo) != int(hi):
            raise ValueError("every rank must hold the same (padded) number of pixels")
        _SIZES_CHECKED.add((rows_pad, G, id(group)))
    cols_pad = rows_pad * G
    C = num_classes
    f32, i32 = torch.float32, torch.int32

    ehat_all = stages.empty((cols_pad, dim), torch.bfloat16)
    bits_all = stages.empty((cols_pad,), i32)
    tor_all = stages.empty((cols_pad // 128,), i32)
    w_all = stages.empty((cols_pad,), f32)
    sl = slice(rank * rows_pad, (rank + 1) * rows_pad)
    tl = slice(rank * (rows_pad // 128), (rank + 1) * (rows_pad // 128))
    ehat, bits, tor, w = ehat_all[sl], bits_all[sl], tor_all[tl], w_all[sl]
    inv_norm = stages.empty((rows_pad,), f32)
    perm = stages.empty((max(n, 1),), i32)
    seg = stages.empty((C + 2,), i32)
    gpos = stages.empty((rows_pad,), f32)
    sym = G > 1 and _use_sym_ring(stages, dim)
    # everything that starts at zero, in ONE buffer and one fill: [row sums | class_sum | class_mass | dproto | loss_acc].
    # The first three are also what the ranks sum up after the row-sum pass, as ONE all-reduce (the class sums are
    # not needed before `mid`, so they ride with the row sums instead of paying a collective of their own).
    n_rs = cols_pad if sym else rows_pad
    n_red = n_rs + C * dim + C
    off_loss = (n_red + C * dim + 1) // 2 * 2           # the double sits on an 8-byte boundary
    zbuf = stages.zeros((off_loss + 2,), f32)
    red = zbuf[:n_red]
    stats = zbuf[n_rs:n_red]                            # class_sum | class_mass
    class_sum, class_mass = stats[:C * dim], stats[C * dim:]
    dproto = zbuf[n_red:n_red + C * dim]
    loss_acc = zbuf[off_loss:off_loss + 2].view(torch.float64)
    loss = stages.empty((), f32)

    stages.prepare(xc, layout, batch, ppi, dim, cls_c, C, eps, ehat, bits, inv_norm, perm, tor, seg, class_sum,
                   class_mass)
    if G > 1:
        _all_gather_many([(ehat_all, ehat), (bits_all, bits), (tor_all, tor)], group)
    unit_off = stages.empty((rows_pad // stages.unit_rows(dim) + 1,), i32)
    stages.plan(tor, rows_pad, tor_all, cols_pad, dim, unit_off)
    if sym:
        # The global similarity matrix is symmetric: every unordered tile pair is computed ONCE.  Every rank numbers the
        # live pairs of the gathered array identically and computes an equal contiguous share of that list (split by
        # count, not by pixel block: the ranks finish together whatever their own images' class mix); a rank's pass
        # leaves row- and column-sum contributions for ANY pixel in rowsum_all, and one all-reduce of N_g floats
        # completes them.  Half the products and exponentials of the ordered pass for 4 bytes per global pixel.
        rowsum_all = zbuf[:cols_pad]
        unit_off_sym = stages.empty((cols_pad // 128 + 1,), i32)
        if hasattr(stages, "rowsum_sym_share") and os.environ.get("MUSCLE_ROWSUM_RING", "0") != "1":
            stages.plan_sym_ring(tor_all, cols_pad, tor_all, cols_pad, 0, 1, unit_off_sym)
            stages.rowsum_sym_share(ehat_all, bits_all, tor_all, cols_pad, unit_off_sym, rank, G, dim, temperature,
                                    rowsum_all)
        else:       # the split by pixel block (ring ownership; kept for A/B measurements: MUSCLE_ROWSUM_RING=1)
            stages.plan_sym_ring(tor, rows_pad, tor_all, cols_pad, rank, G, unit_off_sym)
            stages.rowsum_sym_ring(ehat, bits, tor, rows_pad, ehat_all, bits_all, tor_all, cols_pad, unit_off_sym, rank,
                                   G, dim, temperature, rowsum_all)
        dist.all_reduce(red, group=group)
        rowsum = rowsum_all[sl]
    else:
        rowsum = zbuf[:rows_pad]
        stages.rowsum(ehat, bits, tor, rows_pad, ehat_all, bits_all, tor_all, cols_pad, unit_off, dim, temperature,
                      rowsum)
        if G > 1:
            dist.all_reduce(stats, group=group)
    stages.mid(ehat, seg, rows_pad, dim, C, class_sum, class_mass, rowsum, temperature, w, gpos, loss_acc, dproto)
    if G > 1:
        # everything the gradient half needs from the other ranks is exchanged here, in two launches: the loss
        # and prototype-gradient sums and every pixel's 1/rowsum (N_g floats) — the backward then runs without a
        # collective, and no rank waits for another between the two pair passes' halves
        dist.all_reduce(loss_acc, group=group)            # (float64: cannot share a coalesced call with the fp32 sums)
        if need_grad:
            dist.all_reduce(dproto, group=group)
            _all_gather_into(w_all, w, group)
    part0 = stages.empty((1,), f32)
    stages.final(ehat, bits, inv_norm, perm, part0, unit_off, rows_pad, cols_pad, batch, ppi, dim, C, class_sum,
                 class_mass, gpos, dproto, loss_acc, temperature, layout, None, loss)
    state = None
    if need_grad:
        state = dict(G=G, group=group, xc_shape=tuple(xc.shape), layout=layout, batch=batch, ppi=ppi, dim=dim, C=C,
                     temperature=temperature, rows_pad=rows_pad, cols_pad=cols_pad, ehat=ehat, bits=bits, tor=tor,
                     w=w, ehat_all=ehat_all, bits_all=bits_all, tor_all=tor_all, w_all=w_all, inv_norm=inv_norm,
                     perm=perm, unit_off=unit_off, class_sum=class_sum, class_mass=class_mass, gpos=gpos,
                     dproto=dproto, loss_acc=loss_acc)
    return loss, state


def sharded_bwd(stages, st, grad_scale=None):
    """The gradient half: the gradient pass and the final kernel with the upstream gradient `grad_scale` (0-dim
    tensor or None) multiplied in (the exchange of `w` and of the prototype gradient happened at the end of the
    forward half).  -> dx in the layout of the forward input."""
    f32 = torch.float32
    dim, rows_pad, cols_pad = st["dim"], st["rows_pad"], st["cols_pad"]
    part = stages.empty((stages.part_slots(), rows_pad, dim), f32)
    stages.grad(st["ehat"], st["bits"], st["tor"], st["w"], rows_pad, st["ehat_all"], st["bits_all"], st["tor_all"],
                st["w_all"], cols_pad, st["unit_off"], dim, st["temperature"], part)
    dx = stages.empty(st["xc_shape"], f32)
    stages.final(st["ehat"], st["bits"], st["inv_norm"], st["perm"], part, st["unit_off"], rows_pad, cols_pad,
                 st["batch"], st["ppi"], dim, st["C"], st["class_sum"], st["class_mass"], st["gpos"], st["dproto"],
                 st["loss_acc"], st["temperature"], st["layout"], dx, None, grad_scale)
    return dx


def sharded_fwd_bwd(stages, x, cls, temperature, num_classes, eps, need_grad, group=None):
    """Forward and (if `need_grad`) gradient in one go.  -> (loss 0-dim tensor, dx or None)."""
    loss, st = sharded_fwd(stages, x, cls, temperature, num_classes, eps, need_grad, group)
    return loss, (sharded_bwd(stages, st) if st is not None else None)


class _ShardedDenseInfoNCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cls, temperature, num_classes, eps, group):
        if not x.is_cuda:
            raise RuntimeError("muscle_b200: `x` must be a CUDA tensor (there is no CPU path)")
        if x.dtype != torch.float32:
            raise TypeError("embeddings must be float32")
        stages = CudaStages(x.device)
        loss, st = sharded_fwd(stages, x, cls, temperature, num_classes, eps, x.requires_grad, group)
        ctx.state, ctx.stages = st, stages
        ctx.x_shape = x.shape
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        st = ctx.state
        if st is None:
            return None, None, None, None, None, None
        g = grad_out.detach()
        if g.dtype != torch.float32 or not g.is_contiguous():
            g = g.to(dtype=torch.float32).contiguous()
        dx = sharded_bwd(ctx.stages, st, g)
        ctx.state = None
        return dx.reshape(ctx.x_shape), None, None, None, None, None


def dense_infonce_sharded(x, cls, temperature=0.1, num_classes=21, eps=1e-12, group=None):
    """`muscle_b200.dense_infonce` over the union of all ranks' pixels (cross-GPU negatives).
    Each rank passes its own images; returns the global-mean loss with grad to the local `x`."""
    if num_classes > _lib.MAX_CLASSES:
        raise ValueError("num_classes must be <= %d" % _lib.MAX_CLASSES)
    return _ShardedDenseInfoNCE.apply(x, cls, temperature, num_classes, eps, group)
