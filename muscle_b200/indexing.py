"""the IRN random walk of the reference's ``src/indexing.py`` (:116-150) — same function names and
arguments.  The element-wise half (power, column normalisation; edge masking) runs in libmuscle_b200.so (csrc/irn.cu);
the matrix squarings and the final product are plain dense fp32 GEMMs and go to the GEMM library (``torch.matmul``),
as the reference's result is defined in fp32.  Inference-time refinement: no gradients."""
from __future__ import annotations

import torch

from ._call import call, require_cuda_f32


def to_transition_matrix(affinity_dense, beta, times):
    """src/indexing.py:116-123: ``A ** beta``, columns normalised to sum 1, then ``times`` squarings."""
    require_cuda_f32(affinity_dense, "affinity_dense")
    if affinity_dense.dim() != 2 or affinity_dense.shape[0] != affinity_dense.shape[1]:
        raise ValueError("affinity_dense must be a square matrix")
    a = affinity_dense.detach().contiguous()
    n = a.shape[0]
    colsum = torch.empty(n, dtype=torch.float32, device=a.device)
    trans = torch.empty_like(a)
    call("muscle_irn_transition_init", a.device, a.data_ptr(), n, float(beta), colsum.data_ptr(), trans.data_ptr())
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False          # the reference result is an fp32 product
    try:
        for _ in range(int(times)):
            trans = torch.matmul(trans, trans)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    return trans


def random_walk(x, edge, trans_mat):
    """The tail of ``propagate_to_edge`` (src/indexing.py:144-148): ``(x * (1 - edge)).view(-1, h*w) @ trans_mat``
    -> [C, 1, h, w]."""
    require_cuda_f32(x, "x")
    h, w = x.shape[-2:]
    xf = x.detach().reshape(-1, h * w).contiguous()
    e = edge.detach().reshape(-1).contiguous().float()
    y = torch.empty_like(xf)
    call("muscle_irn_mask_edge", xf.device, xf.data_ptr(), e.data_ptr(), xf.shape[0], h * w, y.data_ptr())
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        rw = torch.matmul(y, trans_mat)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    return rw.reshape(rw.shape[0], 1, h, w)
