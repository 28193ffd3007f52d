"""Stage-by-stage comparison of the fused pre-net with the oracle (bf16 rounding at the same points)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from oracle import uva_oracle as O
from unified_video_action_b200 import DiffActLoss

dev = torch.device("cuda", 0)
torch.manual_seed(11)
mod = DiffActLoss(2, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None)
Pp = {n: p.detach().clone() for n, p in mod.named_parameters() if not n.startswith("net.")}
mod = mod.to(dev)
B = 2
g = torch.Generator().manual_seed(12)
z = torch.randn(B, 1024, 768, generator=g)
with torch.no_grad():
    out = mod.prenet(z.to(dev))
head = mod.net.head()
ws = head._ws[2]
Z, BT, PX = 768, B * 4, B * 4 * 256
off = 0
def take(n):
    global off
    r = off; off += (n + 1023) // 1024 * 1024; return r
o_zb = take(PX * Z * 2); o_pool = take(BT * 16 * Z * 4); o_pb = take(BT * 16 * Z * 2); o_acc = take(BT * Z * 4)
o_h1 = take(BT * Z * 2); o_f2 = take(BT * Z * 4); o_itp = take(B * 16 * Z * 2); o_r1 = take(B * 16 * Z * 2)
pooled = ws[o_pool:o_pool + BT * 16 * Z * 4].view(torch.float32).reshape(BT, 16, Z).cpu()
acc1 = ws[o_acc:o_acc + BT * Z * 4].view(torch.float32).reshape(BT, Z).cpu()
f2 = ws[o_f2:o_f2 + BT * Z * 4].view(torch.float32).reshape(BT, Z).cpu()
r = O._bf16r
x = z.reshape(BT, 256, Z).reshape(BT, 16, 16, Z).permute(0, 3, 1, 2)
conv = F.conv2d(r(x), r(Pp["conv.0.weight"]), Pp["conv.0.bias"], padding=1)
pool_ref = F.adaptive_avg_pool2d(F.relu(conv), (4, 4))  # [BT, Z, 4, 4]
pool_ref_r = pool_ref.permute(0, 2, 3, 1).reshape(BT, 16, Z)
l2 = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
print("pooled vs oracle:", l2(pooled, pool_ref_r))
fc0 = F.linear(r(pool_ref.reshape(BT, -1)), r(Pp["fc.0.weight"]))
print("fc.0 accumulator vs oracle:", l2(acc1, fc0))
h1 = F.relu(fc0 + Pp["fc.0.bias"])
f2r = F.linear(r(h1), r(Pp["fc.2.weight"]), Pp["fc.2.bias"])
print("fc.2 vs oracle:", l2(f2, f2r))
print("out vs oracle:", l2(out.cpu(), O.prenet_conv_fc(Pp, z, emulate_bf16=True)))
