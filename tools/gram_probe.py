"""GPU probe: run gridb200_gram on one random case and compare with numpy (debugging aid, not a test)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from grid_b200._lib import call, ptr

vmax, n, k, impl = (int(x) for x in sys.argv[1:5])
rng = np.random.default_rng(vmax)
a = rng.choice([-vmax, vmax], size=(n, k)).astype(np.int64)
t = torch.from_numpy(np.ascontiguousarray(a.astype(np.float16).reshape(n, k // 64, 64).transpose(1, 0, 2))).cuda()
dot = torch.full((n, n + (n % 2)), -7, dtype=torch.int64, device="cuda")
print("launch", vmax, n, k, impl, flush=True)
call("gridb200_gram", ptr(t), n, n, ptr(t), n, n, k, vmax, ptr(dot), dot.stride(0), 0, impl, torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
ok = (dot.cpu().numpy()[:, :n] == a @ a.T).all()
print("done ok=", ok, flush=True)
