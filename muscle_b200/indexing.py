"""the IRN random walk of the reference's ``src/indexing.py`` (:116-150) — same function names and
arguments.  Everything runs in libmuscle_b200.so (csrc/irn.cu): the element-wise half (power, column normalisation; edge
masking) and the matrix squarings / the final product, which are fp32-accurate products on the tensor pipe (every value
split into three bf16 parts, six part products accumulated in fp32 — the reference's result is defined in fp32, and a
plain bf16 squaring loses the small transition probabilities after three squarings).  ``MUSCLE_IRN_TC=0`` sends the
products to the fp32 GEMM library instead (A/B measurements).  Inference-time refinement: no gradients."""
from __future__ import annotations

import os

import torch

from . import _lib
from ._call import call, require_cuda_f32


def _pad128(v):
    return (int(v) + 127) // 128 * 128


def _tc():
    return os.environ.get("MUSCLE_IRN_TC", "1") != "0"


def _parts(rows_pad, cols_pad, dev):
    nbytes = int(_lib.load().muscle_irn_split3_bytes(rows_pad, cols_pad))
    return torch.empty(nbytes, dtype=torch.uint8, device=dev)


def _split3(x, rows_pad, cols_pad):
    parts = _parts(rows_pad, cols_pad, x.device)
    call("muscle_irn_split3", x.device, x.data_ptr(), x.shape[0], x.shape[1], rows_pad, cols_pad, parts.data_ptr())
    return parts


def _library_matmul(a, b):
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False          # the reference result is an fp32 product
    try:
        return torch.matmul(a, b)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def to_transition_matrix(affinity_dense, beta, times):
    """src/indexing.py:116-123: ``A ** beta``, columns normalised to sum 1, then ``times`` squarings."""
    require_cuda_f32(affinity_dense, "affinity_dense")
    if affinity_dense.dim() != 2 or affinity_dense.shape[0] != affinity_dense.shape[1]:
        raise ValueError("affinity_dense must be a square matrix")
    a = affinity_dense.detach().contiguous()
    n = a.shape[0]
    dev = a.device
    colsum = torch.empty(n, dtype=torch.float32, device=dev)
    trans = torch.empty_like(a)
    call("muscle_irn_transition_init", dev, a.data_ptr(), n, float(beta), colsum.data_ptr(), trans.data_ptr())
    times = int(times)
    if times <= 0:
        return trans
    if not _tc():
        for _ in range(times):
            trans = _library_matmul(trans, trans)
        return trans
    npad = _pad128(n)
    cur = _split3(trans, npad, npad)
    nxt = _parts(npad, npad, dev) if times > 1 else None
    out = torch.empty_like(a)
    for t in range(times):
        last = t == times - 1
        call("muscle_irn_matmul_split3", dev, cur.data_ptr(), cur.data_ptr(), npad, npad, npad, n, n,
             out.data_ptr() if last else None, None if last else nxt.data_ptr())
        cur, nxt = nxt, cur
    return out


def random_walk(x, edge, trans_mat):
    """The tail of ``propagate_to_edge`` (src/indexing.py:144-148): ``(x * (1 - edge)).view(-1, h*w) @ trans_mat``
    -> [C, 1, h, w]."""
    require_cuda_f32(x, "x")
    require_cuda_f32(trans_mat, "trans_mat")
    h, w = x.shape[-2:]
    xf = x.detach().reshape(-1, h * w).contiguous()
    e = edge.detach().reshape(-1).contiguous().float()
    y = torch.empty_like(xf)
    call("muscle_irn_mask_edge", xf.device, xf.data_ptr(), e.data_ptr(), xf.shape[0], h * w, y.data_ptr())
    if not _tc():
        rw = _library_matmul(y, trans_mat)
    else:
        planes, n = y.shape
        mpad, npad = _pad128(planes), _pad128(n)
        ya = _split3(y, mpad, npad)
        tb = _split3(trans_mat.detach().contiguous(), npad, npad)
        rw = torch.empty_like(y)
        call("muscle_irn_matmul_split3", y.device, ya.data_ptr(), tb.data_ptr(), mpad, npad, npad, planes, n, rw.data_ptr(), None)
    return rw.reshape(rw.shape[0], 1, h, w)
