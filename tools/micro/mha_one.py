import sys, os, torch
sys.path.insert(0, os.getcwd())
from kinet_b200.functions.small_mha import small_mha
dev = torch.device("cuda:0")
N, H, D, L = 1, 8, 36, 198
qk = torch.randn(N, L, 2 * H * D, device=dev, dtype=torch.bfloat16)
v = torch.randn(N, L, H * D, device=dev, dtype=torch.bfloat16)
q = qk[..., :H * D].unflatten(-1, (H, D)); k = qk[..., H * D:].unflatten(-1, (H, D)); v4 = v.unflatten(-1, (H, D))
for _ in range(5):
    o = small_mha(q, k, v4)
torch.cuda.synchronize()
