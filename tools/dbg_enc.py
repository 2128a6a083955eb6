import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import tacotron2_oracle as O
from tacotron2_b200 import synth
from tacotron2_b200.model import Tacotron2
a = synth.VANILLA
B, L = int(sys.argv[1]), int(sys.argv[2])
sd = synth.synth_state_dict(a)
m = Tacotron2(**a.kwargs()).cuda(); m.load_state_dict(sd); m.train()
batch = synth.synth_batch(a, B, L, 8, seed=21)
keep = synth.synth_masks(a, B, L, 4, seed=17, train=True)["encoder"]
g = torch.Generator().manual_seed(2)
dy = torch.randn(B, L, a.encoded_dim, generator=g)
sdg = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v.clone()) for k, v in sd.items()}
want = O.encoder(sdg, batch["chars_idx"], batch["chars_idx_len"], train=True, keep=keep, p=a.dropout)
(want * dy[:, : want.shape[1]]).sum().backward()
got = m.encoder(batch["chars_idx"].cuda(), batch["chars_idx_len"].cuda(), dropout_keep=keep)
(got * dy.cuda()).sum().backward()
print("lens", batch["chars_idx_len"].tolist(), "fwd err", float((got.detach().cpu()[:, :want.shape[1]] - want.detach()).abs().max()))
for k, p in m.named_parameters():
    if k.startswith("encoder."):
        w = sdg[k].grad if sdg[k].grad is not None else torch.zeros_like(sdg[k])
        print(f"{k:45s} err {float((p.grad.cpu()-w).abs().max()):.3e} max {float(w.abs().max()):.3e}")
# ---- intermediate check: dgates (still intact in the workspace) -> d_x3 on the CPU vs the oracle's gradient at the conv-stack output
torch.cuda.synchronize()
sdg2 = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v.clone()) for k, v in sd.items()}
conv = O.encoder_convs(sdg2, batch["chars_idx"], True, keep, a.dropout)
conv.retain_grad()
out = O.bilstm_packed(sdg2, conv, batch["chars_idx_len"])
(out * dy).sum().backward()
off = 64 * 2048 * 16 + 2 * 2048 * 4 + 512
dg = m._ws[off: off + B * L * 2048 * 4].view(torch.float32).view(B * L, 2048).cpu()
Wcat = torch.cat([sd["encoder.lstm.weight_ih_l0"], sd["encoder.lstm.weight_ih_l0_reverse"]], 0)
dx3 = (dg.double() @ Wcat.double()).float().view(B, L, 512)
print("d_x3 (from GPU dgates, CPU matmul) vs oracle:", float((dx3 - conv.grad).abs().max()), "max", float(conv.grad.abs().max()))
lens = batch["chars_idx_len"]
for b in range(B):
    e = (dx3[b] - conv.grad[b]).abs().max(dim=1).values
    bad = (e > 1e-3).nonzero().flatten().tolist()
    print("b", b, "len", int(lens[b]), "bad positions", bad[:20], "n", len(bad))
if os.environ.get("T2B200_DEBUG_ENC_STOP"):
    def r256(x): return (x + 255) // 256 * 256
    Lp = (L + 7) // 8 * 8
    o = 0
    for nbytes in (64 * 2048 * 16 + 2 * 2048 * 4 + 512, B * L * 2048 * 4, 2 * B * L * 2048 * 2, 2 * 2048 * B * Lp * 2, 2 * 2 * 256 * B * Lp * 2,
                   5 * 2 * 512 * B * Lp * 2, 2048 * 512 * 4, 2048 * 4):
        o += r256(nbytes)
    dyA = m._ws[o: o + B * L * 512 * 4].view(torch.float32).view(B, L, 512).cpu()
    print("GPU d_x3 (conv_tc k=1) vs oracle:", float((dyA - conv.grad).abs().max()))
    for b in range(B):
        e = (dyA[b] - conv.grad[b]).abs().max(dim=1).values
        bad = (e > 1e-3).nonzero().flatten().tolist()
        print("b", b, "len", int(lens[b]), "bad positions", bad[:40], "n", len(bad), "maxerr", float(e.max()))
if os.environ.get("T2B200_DEBUG_ENC_STOP") == "2":
    import torch.nn.functional as F
    sd3 = {k: v.clone() for k, v in sd.items()}
    x = F.embedding(batch["chars_idx"], sd3["encoder.embedding.weight"], padding_idx=0).transpose(1, 2)
    zs = []
    for li, i in enumerate((0, 4, 8)):
        z = F.conv1d(x, sd3[f"encoder.convolutions.{i}.weight"], sd3[f"encoder.convolutions.{i}.bias"], padding=2)
        z = z.detach().requires_grad_(True) if li == 2 else z
        zs.append(z)
        y, mean, var = O.batchnorm(z, sd3, f"encoder.convolutions.{i + 1}.", True)
        x = O._drop(torch.relu(y), keep[li].transpose(1, 2), a.dropout)
    out = O.bilstm_packed(sd3, x.transpose(1, 2), batch["chars_idx_len"])
    (out * dy).sum().backward()
    dz_or = zs[2].grad.transpose(1, 2)  # (B,L,E)
    o += r256(B * L * 512 * 4) * 2   # skip dyA, dyB
    dzf = m._ws[o: o + B * L * 512 * 4].view(torch.float32).view(B, L, 512).cpu()
    print("GPU dz (layer 2) vs oracle:", float((dzf - dz_or).abs().max()), "max", float(dz_or.abs().max()))
    tp = m._dbg_tape
    to = 4 * r256(2 * B * L * 512 * 2) + 2 * (r256(B * L * 512 * 4) + 2 * r256(512 * 4))
    z2 = tp[to: to + B * L * 512 * 4].view(torch.float32).view(B, L, 512).cpu()
    to += r256(B * L * 512 * 4)
    mean2 = tp[to: to + 2048].view(torch.float32).cpu(); to += r256(2048)
    rstd2 = tp[to: to + 2048].view(torch.float32).cpu()
    zt = zs[2].detach().transpose(1, 2)
    print("tape z2 err", float((z2 - zt).abs().max()), "mean err", float((mean2 - zt.reshape(-1, 512).mean(0)).abs().max()),
          "rstd err", float((rstd2 - torch.rsqrt(zt.reshape(-1, 512).var(0, unbiased=False) + 1e-5)).abs().max()))
    e = (dzf - dz_or).abs()
    print("rows with error > 1e-3:", (e.max(dim=2).values > 1e-3).nonzero().tolist()[:30])
    print("channels with error > 1e-3:", (e.amax(dim=(0, 1)) > 1e-3).nonzero().flatten().tolist()[:40])
if os.environ.get("T2B200_DEBUG_ENC_STOP") == "2":
    ch = (e.amax(dim=(0, 1)) > 1e-3).nonzero().flatten().tolist()
    for cch in ch[:3]:
        zc = zt[:, :, cch].reshape(-1)
        gam, bet = float(sd["encoder.convolutions.9.weight"][cch]), float(sd["encoder.convolutions.9.bias"][cch])
        print("channel", cch, "var", float(zc.var(unbiased=False)), "mean", float(zc.mean()), "gamma", gam, "beta", bet,
              "tape mean", float(mean2[cch]), "tape rstd", float(rstd2[cch]))
        a_t = (z2[:, :, cch] - mean2[cch]) * rstd2[cch] * gam + bet
        a_o = (zt[:, :, cch] - zc.mean()) * torch.rsqrt(zc.var(unbiased=False) + 1e-5) * gam + bet
        print("  relu mask mismatches", int(((a_t > 0) != (a_o > 0)).sum()), "min |a|", float(a_o.abs().min()), "a err", float((a_t - a_o).abs().max()))
        print("  dz gpu", dzf[0, :6, cch].tolist()); print("  dz ora", dz_or[0, :6, cch].tolist())
        print("  dy", dyA[0, :6, cch].tolist() if False else "", "keep", keep[2][0, :6, cch].tolist())
