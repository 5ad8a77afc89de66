import os, sys, torch
sys.path.insert(0, os.getcwd())
import exorl_b200 as xb
from exorl_b200 import knn as xknn
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
E, L = 200, 1000
start = torch.randn(E, 1, 24, generator=g, device=dev)
walk = torch.cumsum(0.05 * torch.randn(E, L, 24, generator=g, device=dev), dim=1)
obs = (start + walk).reshape(-1, 24)
w = torch.randn(512, 24, generator=g, device=dev) / 24 ** 0.5
z = torch.zeros(512, device=dev)
p = xb.icm_get_rep((w, z, torch.ones(512, device=dev), z), obs)
val, idx, cnt, ha, ea, sc = xknn.knn_screen_lists(p, p, 12)
print("lists shape", tuple(cnt.shape), "scalars", sc.view(torch.float32)[:2].tolist(), sc[2].item())
bad = cnt < 0
print("overflow fraction per list:", bad.float().mean(dim=0).tolist())
rows_bad = bad.any(dim=1)
print("rows undecided:", int(rows_bad.sum()))
pos = torch.arange(E * L, device=dev) % L
hist = torch.zeros(10, device=dev)
hist.index_add_(0, (pos[rows_bad] // 100), torch.ones(int(rows_bad.sum()), device=dev))
print("by position in episode (deciles):", hist.tolist())
ep = torch.arange(E * L, device=dev) // L
he = torch.zeros(E, device=dev); he.index_add_(0, ep[rows_bad], torch.ones(int(rows_bad.sum()), device=dev))
print("by episode (first 20):", he[:20].tolist(), "max", he.max().item(), "min", he.min().item())
tile = torch.arange(E * L, device=dev) // 128
ht = torch.zeros(int(tile.max()) + 1, device=dev); ht.index_add_(0, tile[rows_bad], torch.ones(int(rows_bad.sum()), device=dev))
print("tiles with any undecided:", int((ht > 0).sum()), "of", ht.numel(), "full tiles bad:", int((ht == 128).sum()))
good = ~rows_bad
print("cnt stats of decided rows: mean", cnt[good].float().mean().item(), "max", cnt[good].max().item())
# a bad row: look at its true neighbours
r = int(rows_bad.nonzero()[0])
d, i = xb.knn_topk(p.rep[r:r+1], p.rep, 30, mode="exact")
print("row", r, "exact 30-NN dists", [round(x, 3) for x in d[0].tolist()])
print("ha", ha[r].item(), "ea", ea[r].item())
