import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import torch
from trase_b200 import synth, _lib
from helpers import scene_args, cuda_forward, cuda_backward, views
dev = torch.device("cuda:0")
P, W, H, seed = 300000, 1352, 1014, 2
sc = synth.make_scene(P, W, H, seed)
a = scene_args(sc, 0.0, dev)
up = {k: v.to(dev) for k, v in synth.make_upstream_grads(W, H, seed).items()}
ref_in = {k: v.clone() for k, v in a.items() if torch.is_tensor(v)}
for ppt in (2, 1, 2):
    _lib.set_tuning("fwd_ppt", ppt); _lib.set_tuning("bwd_ppt", ppt)
    Rs = set()
    for it in range(12):
        res = cuda_forward(a)
        v = views(res, a)
        Rs.add((res.num_rendered, int(v["tiles_touched"].sum()), int(res.radii.sum()), int(v["counters"][0]), int(v["counters"][1])))
        cuda_backward(a, res, up["dL_dcolor"], up["dL_dobjects"])
    torch.cuda.synchronize()
    changed = [k for k, v in ref_in.items() if not torch.equal(v, a[k])]
    print("ppt", ppt, "distinct (R, sum tiles, sum radii, cnt0, cnt1):", Rs, "inputs changed:", changed)
