"""Rasterizer-only fwd+bwd at config 2 with per-stage device times (library event pairs)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from trase_b200 import synth, _lib
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from helpers import scene_args, cuda_forward, cuda_backward
dev = torch.device("cuda:0")
P, W, H, seed = 300000, 1352, 1014, 2
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for kv in sys.argv[2:]:
    k, v = kv.split("="); _lib.set_tuning(k, int(v))
sc = synth.make_scene(P, W, H, seed)
a = scene_args(sc, 0.0, dev)
up = {k: v.to(dev) for k, v in synth.make_upstream_grads(W, H, seed).items()}
_lib.load()
for _ in range(3):
    res = cuda_forward(a); cuda_backward(a, res, up["dL_dcolor"], up["dL_dobjects"])
torch.cuda.synchronize()
_lib.profile_enable(True)
for _ in range(reps):
    res = cuda_forward(a); cuda_backward(a, res, up["dL_dcolor"], up["dL_dobjects"])
torch.cuda.synchronize()
prof = _lib.profile_read(); _lib.profile_enable(False)
Rs = set()
for _ in range(5):
    Rs.add(cuda_forward(a).num_rendered)
print("R =", res.num_rendered, Rs, " ".join("%s=%.3f" % (k, v[0] / reps) for k, v in prof.items() if v[1]))
