import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import torch
from trase_b200 import synth
from helpers import scene_args, cuda_forward, cuda_backward
dev = torch.device("cuda:0")
P, W, H = 3000, 100, 75
sc = synth.make_scene(P, W, H, 5)
a = scene_args(sc, 0.3, dev)
up = {k: v.to(dev) for k, v in synth.make_upstream_grads(W, H, 5).items()}
for _ in range(2):
    res = cuda_forward(a)
    g = cuda_backward(a, res, up["dL_dcolor"], up["dL_dobjects"])
torch.cuda.synchronize()
print("ok R", res.num_rendered)
