"""CPU stand-in for `muscle_b200.stages.CudaStages` — TEST INFRASTRUCTURE ONLY.

It restates, in torch-on-CPU, what every staged C-ABI call of include/muscle_b200.h ("Staged
form") is documented to leave in its output buffers, so the host-side orchestration of
`muscle_b200.dist.sharded_fwd_bwd` (buffer slicing, the order and payload of the collectives,
which sums are global) can be exercised with the gloo backend at world_size 2 on a machine
without a GPU.  It is never imported by the product package.
"""
import numpy as np
import torch

TOR_INVALID = np.uint32(0x80000000)


def _u32(t):
    """int32 tensor -> numpy uint32 view (shares memory)."""
    return t.numpy().view(np.uint32)


class CpuStages:
    def __init__(self):
        self.device = torch.device("cpu")

    # -- sizes ---------------------------------------------------------------------------------
    def pad_rows(self, n):
        return (max(int(n), 1) + 255) // 256 * 256

    def part_slots(self):
        return 4

    def unit_rows(self, dim):
        return 128

    def empty(self, shape, dtype):
        # poison instead of garbage: reading a buffer nobody wrote shows up as NaN / huge values
        t = torch.empty(shape, dtype=dtype)
        if dtype.is_floating_point:
            t.fill_(float("nan"))
        else:
            t.fill_(0x5A5A5A5A if dtype == torch.int32 else 0)
        return t

    def zeros(self, shape, dtype):
        return torch.zeros(shape, dtype=dtype)

    # -- stages --------------------------------------------------------------------------------
    def prepare(self, x, layout, batch, ppi, dim, cls, num_classes, eps, ehat, bits, inv_norm, perm, tor, seg,
                class_sum, class_mass):
        n = batch * ppi
        flat = x.reshape(batch, dim, ppi).permute(0, 2, 1).reshape(n, dim) if layout == 1 else x.reshape(n, dim)
        cls = cls.reshape(-1).to(torch.int64)
        valid = (cls >= 0) & (cls < num_classes)
        key = torch.where(valid, cls, torch.full_like(cls, num_classes))
        order = torch.argsort(key, stable=True)           # sorted row r holds source pixel order[r]
        perm[order] = torch.arange(n, dtype=torch.int32)
        norm = flat.norm(dim=1).clamp_min(eps)
        e = (flat / norm[:, None]).to(torch.bfloat16)
        ehat.zero_()
        ehat[:n] = e[order]
        b = torch.where(valid, torch.ones_like(cls) << cls.clamp_min(0), torch.zeros_like(cls)).to(torch.int32)
        bits.zero_()
        bits[:n] = b[order]
        inv_norm.zero_()
        inv_norm[:n] = (1.0 / norm)[order]
        bu = _u32(bits).reshape(-1, 128)
        t = np.bitwise_or.reduce(bu, axis=1)
        t = np.where((bu == 0).any(axis=1), t | TOR_INVALID, t).astype(np.uint32)
        _u32(tor)[:] = t
        counts = torch.bincount(key, minlength=num_classes + 1)
        starts = torch.cumsum(counts, 0) - counts
        seg[:num_classes + 1] = starts.to(torch.int32)
        seg[num_classes + 1] = n
        ef = ehat[:n].float()
        sk = key[order]
        for c in range(num_classes):
            m = sk == c
            class_sum[c * dim:(c + 1) * dim] += ef[m].sum(0)
            class_mass[c] += float(m.sum())

    def plan(self, tor_rows, rows_pad, tor_cols, cols_pad, dim, unit_off):
        unit_off.zero_()   # the stand-in's pair passes do not partition work

    @staticmethod
    def _pair_weights(ehat_rows, bits_rows, ehat_cols, bits_cols, temperature):
        s = ehat_rows.float() @ ehat_cols.float().t()
        br, bc = bits_rows.to(torch.int64), bits_cols.to(torch.int64)
        mask = (br[:, None] != bc[None, :]) & (br[:, None] != 0) & (bc[None, :] != 0)
        return torch.exp((s - 1.0) / temperature) * mask

    def rowsum(self, ehat_rows, bits_rows, tor_rows, rows_pad, ehat_cols, bits_cols, tor_cols, cols_pad, unit_off,
               dim, temperature, rowsum):
        rowsum += self._pair_weights(ehat_rows, bits_rows, ehat_cols, bits_cols, temperature).sum(1)

    # -- ring-symmetric row sums (include/muscle_b200.h: muscle_infonce_rowsum_sym_ring) --------------
    def sym_dims(self):
        return (64, 128)

    @staticmethod
    def sym_owned(r_loc, j, tiles, g, G):
        bj, jl = divmod(j, tiles)
        o = (bj - g) % G
        if o == 0:
            return jl >= r_loc
        if 2 * o < G:
            return True
        if 2 * o == G:
            return ((r_loc + jl) & 1) == (0 if g < bj else 1)
        return False

    def plan_sym_ring(self, tor_rows, rows_pad, tor_cols, cols_pad, rank, world, unit_off_sym):
        unit_off_sym.zero_()

    def rowsum_sym_ring(self, ehat_rows, bits_rows, tor_rows, rows_pad, ehat_cols, bits_cols, tor_cols, cols_pad,
                        unit_off_sym, rank, world, dim, temperature, rowsum_all):
        """every unordered tile pair on the rank that owns it: row sums into this rank's slice of rowsum_all, column
        sums into the column pixels' entries (off-diagonal tiles only)"""
        p = self._pair_weights(ehat_rows, bits_rows, ehat_cols, bits_cols, temperature)
        tiles = rows_pad // 128
        for r in range(tiles):
            for j in range(cols_pad // 128):
                if not self.sym_owned(r, j, tiles, rank, world):
                    continue
                blk = p[r * 128:(r + 1) * 128, j * 128:(j + 1) * 128]
                rowsum_all[rank * rows_pad + r * 128:rank * rows_pad + (r + 1) * 128] += blk.sum(1)
                if j != rank * tiles + r:
                    rowsum_all[j * 128:(j + 1) * 128] += blk.sum(0)

    def rowsum_sym_share(self, ehat_all, bits_all, tor_all, cols_pad, unit_off_sym, share, n_shares, dim, temperature,
                         rowsum_all):
        """the live unordered tile pairs of the whole gathered array, numbered row-major from the diagonal on
        (include/muscle_b200.h: muscle_infonce_rowsum_sym_share); this rank computes one contiguous share"""
        tiles = cols_pad // 128
        t = _u32(tor_all).astype(np.int64) & 0x7fffffff
        units = []
        for r in range(tiles):
            for j in range(r, tiles):
                dead = t[r] == 0 or t[j] == 0 or (t[r] == t[j] and (t[r] & (t[r] - 1)) == 0)
                if not dead:
                    units.append((r, j))
        lo, hi = len(units) * share // n_shares, len(units) * (share + 1) // n_shares
        for r, j in units[lo:hi]:
            blk = self._pair_weights(ehat_all[r * 128:(r + 1) * 128], bits_all[r * 128:(r + 1) * 128],
                                     ehat_all[j * 128:(j + 1) * 128], bits_all[j * 128:(j + 1) * 128], temperature)
            rowsum_all[r * 128:(r + 1) * 128] += blk.sum(1)
            if j != r:
                rowsum_all[j * 128:(j + 1) * 128] += blk.sum(0)

    def mid(self, ehat, seg, rows_pad, dim, num_classes, class_sum, class_mass, rowsum, temperature, w, gpos,
            loss_acc, dproto):
        w.zero_()
        gpos.zero_()
        it = 1.0 / temperature
        for c in range(num_classes):
            lo, hi = int(seg[c]), int(seg[c + 1])
            if hi <= lo:
                continue
            e = ehat[lo:hi].float()
            proto = class_sum[c * dim:(c + 1) * dim] / class_mass[c]
            pos = (e @ proto) * it
            ep = torch.exp(pos - it)
            rs = rowsum[lo:hi] + ep
            rowsum[lo:hi] = rs
            w[lo:hi] = 1.0 / rs
            g = (ep / rs - 1.0) * it
            gpos[lo:hi] = g
            loss_acc += (it + torch.log(rs) - pos).double().sum()
            dproto[c * dim:(c + 1) * dim] += (g[:, None] * e).sum(0)

    def grad(self, ehat_rows, bits_rows, tor_rows, w_rows, rows_pad, ehat_cols, bits_cols, tor_cols, w_cols,
             cols_pad, unit_off, dim, temperature, part):
        p = self._pair_weights(ehat_rows, bits_rows, ehat_cols, bits_cols, temperature)
        t = p * (w_rows[:, None] + w_cols[None, :])
        part.zero_()
        part[0] = t @ ehat_cols.float()

    def final(self, ehat, bits, inv_norm, perm, part, unit_off, rows_pad, cols_pad, batch, ppi, dim, num_classes,
              class_sum, class_mass, gpos, dproto, loss_acc, temperature, layout, dx, loss_out, grad_scale=None):
        n_valid = class_mass.sum()
        if loss_out is not None:
            loss_out.copy_((loss_acc[0] / n_valid.double()).float())
        if dx is None:
            return
        n = batch * ppi
        q = perm[:n].to(torch.int64)
        bq = bits[q].to(torch.int64)
        ok = bq != 0
        c = torch.where(ok, torch.log2(bq.clamp_min(1).double()).round().to(torch.int64), torch.zeros_like(bq))
        e = ehat[q].float()
        cs = class_sum.reshape(num_classes, dim)
        dp = dproto.reshape(num_classes, dim)
        inv_mass = torch.where(class_mass > 0, 1.0 / class_mass, torch.zeros_like(class_mass))[c]
        g = (part.sum(0)[q] / temperature + (gpos[q] * inv_mass)[:, None] * cs[c] + inv_mass[:, None] * dp[c])
        g = g / n_valid
        g = inv_norm[q][:, None] * (g - e * (g * e).sum(1, keepdim=True))
        g = g * ok[:, None]
        if grad_scale is not None:
            g = g * grad_scale.reshape(()).float()
        if layout == 1:
            dx.copy_(g.reshape(batch, ppi, dim).permute(0, 2, 1).reshape(dx.shape))
        else:
            dx.copy_(g.reshape(dx.shape))
