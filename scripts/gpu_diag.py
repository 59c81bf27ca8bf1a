"""First-contact diagnostics on a B200: tcgen05 building blocks, then the dense InfoNCE against
the CPU oracle.  Prints detail (not just pass/fail) so one remote run localises a fault."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import muscle_b200 as M
from oracle import muscle_oracle as O

torch.manual_seed(0)
dev = torch.device("cuda:0")
print(torch.cuda.get_device_name(0))

def tile_check(d):
    a = (torch.randn(128, d) / d ** 0.5).to(torch.bfloat16)
    b = (torch.randn(128, d) / d ** 0.5).to(torch.bfloat16)
    s, g = M.ops.debug_tile(a.to(dev), b.to(dev))
    torch.cuda.synchronize()
    s_ref = a.float() @ b.float().t()
    g_ref = s_ref.to(torch.bfloat16).float() @ b.float()
    es = (s.cpu() - s_ref).abs().max().item()
    eg = (g.cpu() - g_ref).abs().max().item()
    print("debug_tile D=%d  max|S-ref|=%.3e (|S|max %.3f)  max|G-ref|=%.3e (|G|max %.3f)"
          % (d, es, s_ref.abs().max().item(), eg, g_ref.abs().max().item()))
    if es > 1e-2:
        bad = ((s.cpu() - s_ref).abs() > 1e-2)
        print("  S bad rows:", bad.any(1).nonzero().flatten()[:16].tolist(), " bad cols:", bad.any(0).nonzero().flatten()[:16].tolist())
        print("  S[0,:8]  got", s[0, :8].tolist(), "\n           ref", s_ref[0, :8].tolist())
    if eg > 2e-2:
        bad = ((g.cpu() - g_ref).abs() > 2e-2)
        print("  G bad rows:", bad.any(1).nonzero().flatten()[:16].tolist(), " bad cols:", bad.any(0).nonzero().flatten()[:16].tolist())
        print("  G[0,:8]  got", g[0, :8].tolist(), "\n           ref", g_ref[0, :8].tolist())
    return es < 1e-2 and eg < 2e-2

ok = True
for d in (64, 128, 256):
    try:
        ok &= tile_check(d)
    except Exception as e:
        print("debug_tile D=%d FAILED: %r" % (d, e)); ok = False

def loss_check(batch, d, h, w, seed, layout="nchw"):
    x, cls, _ = O.synth_embeddings(batch, d, h, w, seed)
    xr = x.clone().requires_grad_()
    flat = xr.permute(0, 2, 3, 1).reshape(-1, d)
    ref = O.dense_infonce_from_features(flat, cls.reshape(-1), 0.1, 21)
    gref, = torch.autograd.grad(ref, xr)
    xg = x.to(dev).requires_grad_()
    if layout == "nchw":
        loss = M.dense_infonce(xg, cls.to(dev), 0.1, 21)
    else:
        xp = x.permute(0, 2, 3, 1).reshape(-1, d).contiguous().to(dev).requires_grad_()
        loss = M.dense_infonce(xp, cls.reshape(-1).to(dev), 0.1, 21)
        xg = xp
        gref = gref.permute(0, 2, 3, 1).reshape(-1, d)
    loss.backward()
    torch.cuda.synchronize()
    g = xg.grad.cpu()
    rel = abs(loss.item() - ref.item()) / abs(ref.item())
    cos = torch.nn.functional.cosine_similarity(g.flatten(), gref.flatten(), dim=0).item()
    nr = (g.norm() / gref.norm()).item()
    print("dense_infonce B=%d D=%d %dx%d %s: loss %.6f ref %.6f rel %.2e | grad cos %.6f norm ratio %.4f"
          % (batch, d, h, w, layout, loss.item(), ref.item(), rel, cos, nr))
    return rel < 1e-3 and cos > 0.999

for args in [(2, 128, 16, 16, 1, "nchw"), (2, 128, 16, 16, 1, "pm"), (2, 128, 32, 32, 2, "nchw"), (3, 64, 20, 24, 3, "nchw"),
             (2, 256, 32, 32, 4, "nchw"), (8, 128, 56, 56, 5, "nchw")]:
    try:
        ok &= loss_check(*args)
    except Exception as e:
        print("dense_infonce %s FAILED: %r" % (args, e)); ok = False
        break
print("DIAG", "OK" if ok else "FAILED")
