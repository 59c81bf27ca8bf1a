"""The staged form of the dense InfoNCE (include/muscle_b200.h, "Staged form"), as thin tensor
wrappers over the C ABI.  `muscle_b200.dist` composes these with torch.distributed collectives.
The class is the seam the world_size-2 gloo tests replace with a CPU stand-in."""
from __future__ import annotations

import torch

from . import _lib


def _sp(device):
    return torch.cuda.current_stream(device).cuda_stream


class CudaStages:
    """Every method enqueues one C-ABI call on the current stream of `device`."""

    def __init__(self, device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("muscle_b200 stages need a CUDA device (there is no CPU path)")
        self.lib = _lib.load()

    # -- sizes ---------------------------------------------------------------------------------
    def pad_rows(self, n):
        return int(self.lib.muscle_pad_rows(int(n)))

    def part_slots(self):
        return int(self.lib.muscle_infonce_part_slots())

    def unit_rows(self, dim):
        """rows per row block of the pair passes (the granularity of `unit_off`)"""
        return 128

    def empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def zeros(self, shape, dtype):
        return torch.zeros(shape, dtype=dtype, device=self.device)

    # -- stages --------------------------------------------------------------------------------
    def prepare(self, x, layout, batch, ppi, dim, cls, num_classes, eps, ehat, bits, inv_norm, perm, tor, seg,
                class_sum, class_mass):
        n = batch * ppi
        sb = int(self.lib.muscle_infonce_prepare_scratch_bytes(n))
        scratch = torch.empty(sb, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_prepare(
                x.data_ptr(), layout, batch, ppi, dim, cls.data_ptr(), 1 if cls.dtype == torch.int64 else 0,
                num_classes, float(eps), ehat.data_ptr(), bits.data_ptr(), inv_norm.data_ptr(), perm.data_ptr(),
                tor.data_ptr(), seg.data_ptr(), class_sum.data_ptr(), class_mass.data_ptr(), scratch.data_ptr(), sb,
                _sp(self.device))
        _lib.check(rc, "infonce_prepare")

    def plan(self, tor_rows, rows_pad, tor_cols, cols_pad, dim, unit_off):
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_plan(tor_rows.data_ptr(), rows_pad, tor_cols.data_ptr(), cols_pad, dim,
                                              unit_off.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_plan")

    def rowsum(self, ehat_rows, bits_rows, tor_rows, rows_pad, ehat_cols, bits_cols, tor_cols, cols_pad, unit_off,
               dim, temperature, rowsum):
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_rowsum(
                ehat_rows.data_ptr(), bits_rows.data_ptr(), tor_rows.data_ptr(), rows_pad, ehat_cols.data_ptr(),
                bits_cols.data_ptr(), tor_cols.data_ptr(), cols_pad, unit_off.data_ptr(), dim, float(temperature),
                rowsum.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_rowsum")

    def plan_sym(self, tor, rows_pad, unit_off_sym):
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_plan_sym(tor.data_ptr(), rows_pad, unit_off_sym.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_plan_sym")

    def rowsum_sym(self, ehat, bits, tor, rows_pad, unit_off_sym, dim, temperature, rowsum):
        """rows == columns (single GPU): only the live units with column tile >= row tile"""
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_rowsum_sym(ehat.data_ptr(), bits.data_ptr(), tor.data_ptr(), rows_pad,
                                                    unit_off_sym.data_ptr(), dim, float(temperature),
                                                    rowsum.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_rowsum_sym")

    def plan_sym_ring(self, tor_rows, rows_pad, tor_cols, cols_pad, rank, world, unit_off_sym):
        sb = int(self.lib.muscle_infonce_plan_sym_ring_scratch_bytes(rows_pad))
        scratch = torch.empty(sb, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_plan_sym_ring(tor_rows.data_ptr(), rows_pad, tor_cols.data_ptr(), cols_pad, rank,
                                                       world, scratch.data_ptr(), unit_off_sym.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_plan_sym_ring")

    def rowsum_sym_ring(self, ehat_rows, bits_rows, tor_rows, rows_pad, ehat_cols, bits_cols, tor_cols, cols_pad,
                        unit_off_sym, rank, world, dim, temperature, rowsum_all):
        """symmetric row-sum pass over `world` ranks: every unordered tile pair once, on the rank that owns it;
        rowsum_all [cols_pad] must be all-reduced (sum) afterwards"""
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_rowsum_sym_ring(
                ehat_rows.data_ptr(), bits_rows.data_ptr(), tor_rows.data_ptr(), rows_pad, ehat_cols.data_ptr(),
                bits_cols.data_ptr(), tor_cols.data_ptr(), cols_pad, unit_off_sym.data_ptr(), rank, world, dim,
                float(temperature), rowsum_all.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_rowsum_sym_ring")

    def rowsum_sym_share(self, ehat_all, bits_all, tor_all, cols_pad, unit_off_sym, share, n_shares, dim, temperature,
                         rowsum_all):
        """symmetric row-sum pass over the whole gathered array, split by COUNT: share `share` of `n_shares` equal
        parts of the live unordered tile pairs (plan: plan_sym_ring over the gathered arrays, rank 0 of world 1);
        rowsum_all [cols_pad] must be all-reduced (sum) afterwards"""
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_rowsum_sym_share(
                ehat_all.data_ptr(), bits_all.data_ptr(), tor_all.data_ptr(), cols_pad, unit_off_sym.data_ptr(), share,
                n_shares, dim, float(temperature), rowsum_all.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_rowsum_sym_share")

    def sym_dims(self):
        return (64, 128)

    def mid(self, ehat, seg, rows_pad, dim, num_classes, class_sum, class_mass, rowsum, temperature, w, gpos,
            loss_acc, dproto):
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_mid(
                ehat.data_ptr(), seg.data_ptr(), rows_pad, dim, num_classes, class_sum.data_ptr(),
                class_mass.data_ptr(), rowsum.data_ptr(), float(temperature), w.data_ptr(), gpos.data_ptr(),
                loss_acc.data_ptr(), dproto.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_mid")

    def grad(self, ehat_rows, bits_rows, tor_rows, w_rows, rows_pad, ehat_cols, bits_cols, tor_cols, w_cols,
             cols_pad, unit_off, dim, temperature, part):
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_grad(
                ehat_rows.data_ptr(), bits_rows.data_ptr(), tor_rows.data_ptr(), w_rows.data_ptr(), rows_pad,
                ehat_cols.data_ptr(), bits_cols.data_ptr(), tor_cols.data_ptr(), w_cols.data_ptr(), cols_pad,
                unit_off.data_ptr(), dim, float(temperature), part.data_ptr(), _sp(self.device))
        _lib.check(rc, "infonce_grad")

    def final(self, ehat, bits, inv_norm, perm, part, unit_off, rows_pad, cols_pad, batch, ppi, dim, num_classes,
              class_sum, class_mass, gpos, dproto, loss_acc, temperature, layout, dx, loss_out, grad_scale=None):
        with torch.cuda.device(self.device):
            rc = self.lib.muscle_infonce_final(
                ehat.data_ptr(), bits.data_ptr(), inv_norm.data_ptr(), perm.data_ptr(), part.data_ptr(),
                unit_off.data_ptr(), rows_pad, cols_pad, batch, ppi, dim, num_classes, class_sum.data_ptr(),
                class_mass.data_ptr(), gpos.data_ptr(), dproto.data_ptr(), loss_acc.data_ptr(), float(temperature),
                layout, dx.data_ptr() if dx is not None else None,
                loss_out.data_ptr() if loss_out is not None else None,
                grad_scale.data_ptr() if grad_scale is not None else None, _sp(self.device))
        _lib.check(rc, "infonce_final")
