"""The symmetric row-sum pass (rows == columns: units with column tile >= row tile only, column sums of
every off-diagonal tile on the tensor pipe) against the ordered pass and against the oracle's row sums,
through the staged C ABI."""
import pytest
import torch

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _prepare(st, x, cls, C=21):
    from muscle_b200 import _lib
    dev = st.device
    b, d, h, w = x.shape
    n = b * h * w
    rp = st.pad_rows(n)
    f32, i32 = torch.float32, torch.int32
    ehat = st.empty((rp, d), torch.bfloat16)
    bits, inv_norm = st.empty((rp,), i32), st.empty((rp,), f32)
    tor, perm, seg = st.empty((rp // 128,), i32), st.empty((n,), i32), st.empty((C + 2,), i32)
    stats = st.zeros((C * d + C,), f32)
    st.prepare(x.to(dev), _lib.LAYOUT_NCHW, b, h * w, d, cls.to(dev).reshape(-1), C, 1e-12, ehat, bits, inv_norm, perm,
               tor, seg, stats[:C * d], stats[C * d:])
    return rp, ehat, bits, tor, perm


@pytest.mark.parametrize("batch,d,h,w,temperature,ignore", [
    (2, 128, 32, 32, 0.1, False), (3, 64, 20, 24, 0.1, True), (1, 128, 7, 9, 0.07, False),
    (4, 128, 40, 40, 0.1, True), (2, 128, 19, 23, 0.05, False), (8, 128, 56, 56, 0.1, False),
])
def test_rowsum_sym_matches_ordered_pass_and_oracle(batch, d, h, w, temperature, ignore):
    from muscle_b200.stages import CudaStages
    st = CudaStages(torch.device("cuda", 0))
    x, cls, _ = O.synth_embeddings(batch, d, h, w, seed=300 + batch + d + h)
    if ignore:
        cls = cls.clone()
        cls[0, :3] = -1
        cls[-1, -2:, ::3] = -1
    rp, ehat, bits, tor, perm = _prepare(st, x, cls)
    uo = st.empty((rp // 128 + 1,), torch.int32)
    uo_sym = st.empty((rp // 128 + 1,), torch.int32)
    st.plan(tor, rp, tor, rp, d, uo)
    st.plan_sym(tor, rp, uo_sym)
    rs = st.zeros((rp,), torch.float32)
    rs_sym = st.zeros((rp,), torch.float32)
    st.rowsum(ehat, bits, tor, rp, ehat, bits, tor, rp, uo, d, temperature, rs)
    st.rowsum_sym(ehat, bits, tor, rp, uo_sym, d, temperature, rs_sym)
    torch.cuda.synchronize()
    # every ordered live unit off the diagonal is visited once from either side
    tor_c = tor.cpu().numpy().astype("int64") & 0x7fffffff
    live = 0
    for r in range(len(tor_c)):
        for c in range(r, len(tor_c)):
            a, b_ = int(tor_c[r]), int(tor_c[c])
            dead = a == 0 or b_ == 0 or (a == b_ and (a & (a - 1)) == 0)
            live += 0 if dead else 1
    assert int(uo_sym[-1]) == live
    a, b_ = rs.cpu(), rs_sym.cpu()
    assert torch.isfinite(b_).all()
    torch.testing.assert_close(b_, a, rtol=2e-3, atol=0)
    # oracle: sum over negatives of exp((s - 1)/T) with fp32 unit vectors, in the sorted order
    flat = x.permute(0, 2, 3, 1).reshape(-1, d)
    c = cls.reshape(-1)
    e = O.l2_normalize_pixels(flat.double())
    s = e @ e.t()
    neg = (c[:, None] != c[None, :]) & (c[None, :] >= 0)
    ref = (torch.exp((s - 1) / temperature) * neg).sum(1)
    got = b_[perm.cpu().long()].double()
    ok = c >= 0
    rel = ((got[ok] - ref[ok]).abs() / ref[ok].clamp_min(1e-300))
    assert float(rel.max()) < 2e-2, float(rel.max())          # bf16 operands: per-row sums of e^{s/T}
    assert float(rel.mean()) < 3e-3, float(rel.mean())


@pytest.mark.parametrize("world,batch,d,h,w", [(2, 4, 128, 24, 24), (3, 3, 64, 20, 20), (4, 8, 128, 16, 16), (8, 8, 128, 16, 16)])
def test_ring_symmetric_pass_over_emulated_ranks(world, batch, d, h, w):
    """The multi-rank form on ONE device: every rank's shard is prepared on its own (class-sorted per rank, like
    dist.sharded_fwd), the shards are concatenated the way the all-gather lays them out, and each rank's pass is run
    in turn into one rowsum_all buffer (which is what the all-reduce of the real thing sums up).  Every unordered tile
    pair must then have been counted exactly once: equal to the ordered pass of every rank over all columns."""
    from muscle_b200.stages import CudaStages
    st = CudaStages(torch.device("cuda", 0))
    x, cls, _ = O.synth_embeddings(batch, d, h, w, seed=500 + world)
    cls = cls.clone()
    cls[0, :2] = -1
    lb = batch // world
    shards = [_prepare(st, x[r * lb:(r + 1) * lb], cls[r * lb:(r + 1) * lb]) for r in range(world)]
    rp = shards[0][0]
    assert all(s[0] == rp for s in shards)
    ehat_all = torch.cat([s[1] for s in shards]).contiguous()
    bits_all = torch.cat([s[2] for s in shards]).contiguous()
    tor_all = torch.cat([s[3] for s in shards]).contiguous()
    cp = rp * world
    rs_all = st.zeros((cp,), torch.float32)
    ref_all = st.zeros((cp,), torch.float32)
    total_units = 0
    for r in range(world):
        _, ehat, bits, tor, _ = shards[r]
        uo = st.empty((rp // 128 + 1,), torch.int32)
        st.plan_sym_ring(tor, rp, tor_all, cp, r, world, uo)
        st.rowsum_sym_ring(ehat, bits, tor, rp, ehat_all, bits_all, tor_all, cp, uo, r, world, d, 0.1, rs_all)
        total_units += int(uo[-1])
        uo2 = st.empty((rp // 128 + 1,), torch.int32)
        st.plan(tor, rp, tor_all, cp, d, uo2)
        st.rowsum(ehat, bits, tor, rp, ehat_all, bits_all, tor_all, cp, uo2, d, 0.1, ref_all[r * rp:(r + 1) * rp])
    torch.cuda.synchronize()
    # live unordered pairs (diagonal included) over the concatenated tiles
    tc = tor_all.cpu().numpy().astype("int64") & 0x7fffffff
    live = sum(1 for a in range(len(tc)) for b_ in range(a, len(tc))
               if not (tc[a] == 0 or tc[b_] == 0 or (tc[a] == tc[b_] and (tc[a] & (tc[a] - 1)) == 0)))
    assert total_units == live
    torch.testing.assert_close(rs_all.cpu(), ref_all.cpu(), rtol=2e-3, atol=0)


@pytest.mark.parametrize("world,batch,d,h,w", [(2, 4, 128, 24, 24), (3, 3, 64, 20, 20), (8, 8, 128, 16, 16), (5, 8, 128, 40, 40)])
def test_share_split_symmetric_pass_over_emulated_ranks(world, batch, d, h, w):
    """The split by COUNT (muscle_infonce_rowsum_sym_share): one plan over the gathered array (class-sorted per rank
    block only), every emulated rank computes its contiguous share of the live unordered tile pairs into one
    rowsum_all buffer; the total must equal the ordered pass over all rows x all columns.  (`world` shares need not
    divide the batch: the shards here only fix how the gathered array is laid out.)"""
    from muscle_b200.stages import CudaStages
    st = CudaStages(torch.device("cuda", 0))
    x, cls, _ = O.synth_embeddings(batch, d, h, w, seed=700 + world)
    cls = cls.clone()
    cls[0, :2] = -1
    blocks = 2 if batch % 2 == 0 else batch
    lb = batch // blocks
    shards = [_prepare(st, x[r * lb:(r + 1) * lb], cls[r * lb:(r + 1) * lb]) for r in range(blocks)]
    rp = shards[0][0]
    ehat_all = torch.cat([s[1] for s in shards]).contiguous()
    bits_all = torch.cat([s[2] for s in shards]).contiguous()
    tor_all = torch.cat([s[3] for s in shards]).contiguous()
    cp = rp * blocks
    rs_all = st.zeros((cp,), torch.float32)
    ref_all = st.zeros((cp,), torch.float32)
    uo = st.empty((cp // 128 + 1,), torch.int32)
    st.plan_sym_ring(tor_all, cp, tor_all, cp, 0, 1, uo)
    for share in range(world):
        st.rowsum_sym_share(ehat_all, bits_all, tor_all, cp, uo, share, world, d, 0.1, rs_all)
    uo2 = st.empty((cp // 128 + 1,), torch.int32)
    st.plan(tor_all, cp, tor_all, cp, d, uo2)
    st.rowsum(ehat_all, bits_all, tor_all, cp, ehat_all, bits_all, tor_all, cp, uo2, d, 0.1, ref_all)
    torch.cuda.synchronize()
    tc = tor_all.cpu().numpy().astype("int64") & 0x7fffffff
    live = sum(1 for a in range(len(tc)) for b_ in range(a, len(tc))
               if not (tc[a] == 0 or tc[b_] == 0 or (tc[a] == tc[b_] and (tc[a] & (tc[a] - 1)) == 0)))
    assert int(uo[-1]) == live
    torch.testing.assert_close(rs_all.cpu(), ref_all.cpu(), rtol=2e-3, atol=0)
