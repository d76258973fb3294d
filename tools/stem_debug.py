import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_tutorials_b200.ops import native

N, H, W = [int(v) for v in sys.argv[1:4]]
torch.manual_seed(11)
x = torch.randn(N, 3, H, W, device="cuda")
w = torch.randn(64, 3, 7, 7, device="cuda") / 147 ** 0.5
wp = torch.empty(64, 224, dtype=torch.bfloat16, device="cuda")
native.stem_pack_weight(w.permute(0, 2, 3, 1).contiguous(), wp)
xp = native.stem_pack_input(x)
y, _ = native.stem_fprop(xp, wp, N, H, W, None)
y2 = torch.full_like(y, 7.0)
torch.cuda.synchronize()
ref = torch.nn.functional.conv2d(x.to(torch.bfloat16).float(), w.to(torch.bfloat16).float(), stride=2, padding=3)
err = (y.float() - ref).abs()
err = torch.nan_to_num(err, nan=1e9, posinf=1e9)
bad = err > 0.05
print("shape", tuple(y.shape), "bad frac", bad.float().mean().item())
print("bad by n ", bad.float().mean((1, 2, 3)).tolist())
print("bad by oh", [round(v, 2) for v in bad.float().mean((0, 1, 3)).tolist()])
print("bad by ow", [round(v, 2) for v in bad.float().mean((0, 1, 2)).tolist()])
print("bad by c ", [round(v, 2) for v in bad.float().mean((0, 2, 3)).tolist()][:64])
# --- check the packed image itself
Hp, Wp = H + 6, (W + 9) & ~1
xpv = xp.view(N, Hp, Wp, 4).float()
exp = torch.zeros(N, Hp, Wp, 4, device="cuda")
exp[:, 3:3 + H, 4:4 + W, :3] = x.to(torch.bfloat16).float().permute(0, 2, 3, 1)
print("pack max err", (xpv - exp).abs().max().item(), "numel", xp.numel(), N * Hp * Wp * 4)
# reference from the packed image through the window formulation
wpf = wp.float().view(64, 7, 8, 4)
Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
ref2 = torch.zeros(N, Ho, Wo, 64, device="cuda")
for kh in range(7):
    rows = xpv[:, kh:kh + 2 * Ho:2]                       # [N, Ho, Wp, 4]
    win = rows.reshape(N, Ho, Wp * 4).unfold(2, 32, 8)[:, :, :Wo]   # [N, Ho, Wo, 32]
    ref2 += torch.einsum("nhwk,ck->nhwc", win, wpf[:, kh].reshape(64, 32))
print("window-formulation vs conv2d", (ref2.permute(0, 3, 1, 2) - ref).abs().max().item())
print("kernel vs window-formulation", torch.nan_to_num((y.float() - ref2.permute(0, 3, 1, 2)).abs(), nan=1e9).max().item())
