"""Small end-to-end case for compute-sanitizer (memcheck / racecheck): the config-1-shaped chain (every step-4 kernel, the
symmetric tcgen05 Gram with split-K, k-best merge, dipCN), one multi-tile rectangular Gram block through the dot-matrix
path with the transposed reading, and the same block through the fused Gram + k-best kernel.

    compute-sanitizer --tool memcheck python tools/sanitize_case.py
"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from grid_b200 import engine, synth  # noqa: E402

torch.cuda.set_device(0)
n, b = 96, 1536
p = engine.Params(num_neighbors=20, n_nbr=10)
depth = engine.synth_depth(5, n, b)
counts = torch.from_numpy(synth.read_counts(5, n, [42599])).cuda()
res, idx, d2, cn = engine.run_chain(depth, b, counts, p)
torch.cuda.synchronize()
print("chain ok", int(idx.sum()), float(cn.sum()))

rng = np.random.default_rng(1)
k, kdim = 30, 2048


def operand(rows):
    z = torch.from_numpy(rng.integers(-200, 201, size=(rows, kdim)).astype(np.int16)).cuda()
    return engine.prepare_from_zq(z, kdim, None, 200)


zq, kd, nq_ = operand(520)       # 3 x 3 tiles of 256 with ragged edges
zb, _, nb_ = operand(600)
za, _, na_ = operand(300)
q = engine.knn_init(520, k, "cuda")
engine.knn_block(zq, nq_, 0, za, na_, 5000, kd, 200, q[0], q[1])
bl = engine.knn_init(600, k, "cuda")
engine.knn_block(zb, nb_, 1000, za, na_, 5000, kd, 200, bl[0], bl[1])
t1 = engine.knn_init(600, k, "cuda")
q1 = (q[0].clone(), q[1].clone())
engine.knn_block(zq, nq_, 0, zb, nb_, 1000, kd, 200, q1[0], q1[1], t_lists=t1)
t2 = engine.knn_init(600, k, "cuda")
q2 = (q[0].clone(), q[1].clone())
assert engine.knn_block_fused(zq, nq_, 0, zb, nb_, 1000, kd, 200, q2[0], q2[1], tau_r=bl[0][:, k - 1].contiguous(), t_lists=t2)
torch.cuda.synchronize()
assert torch.equal(q1[0], q2[0]) and torch.equal(q1[1], q2[1])
print("blocks ok")
