"""Small end-to-end pass for compute-sanitizer (tools/sanitize.sh): smoke() of the rasterizer path, a 20k-row
DeformNetwork forward/backward (tcgen05 kernels, CTA pairs, mbarriers), one sort above and one below the 2^20-key
tile switch, the FEATURE-state kernels (smoothed features, feature-only backward)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.smoke()
from trase_b200 import _lib, synth, render as R  # noqa: E402
from trase_b200.deform import DeformNetwork  # noqa: E402
from trase_b200.feature_utils import build_feature_smooth_map  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
net = DeformNetwork(backend="native").to(dev)
M = 20000
x = (torch.rand(M, 3, device=dev) - 0.5) * 3
t = torch.full((M, 1), 0.4, device=dev)
outs = net(x, t)
torch.autograd.backward(list(outs), [torch.randn_like(o) for o in outs])
torch.cuda.synchronize()
print("mlp ok", float(net.linear[0].weight.grad.abs().sum()))

lib = _lib.load()
for n, end_bit in ((1_200_000, 45), (70_000, 39)):
    keys = torch.randint(0, 1 << 40, (n,), device=dev, dtype=torch.int64)
    vals = torch.arange(n, device=dev, dtype=torch.int32)
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    cnt = torch.tensor([n], dtype=torch.int32, device=dev)
    sb = lib.trase_b200_sort_scratch_bytes(n)
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    rc = lib.trase_b200_sort_pairs(keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), cnt.data_ptr(), n,
                                   end_bit, scratch.data_ptr(), sb, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    rc = lib.trase_b200_sort_pairs_static(keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(),
                                          cnt.data_ptr(), n, end_bit, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
torch.cuda.synchronize()
print("sort ok")

P = 5000
sc = synth.make_scene(P, 96, 80, 45)
pc = R.GaussianParams.from_activated(sc, dev)
for p in pc.parameters():
    p.requires_grad_(False)
pc._gaussian_features.requires_grad_(True)
pc.set_feature_smooth_map(build_feature_smooth_map(sc["means3D"], K=16))
z = [torch.zeros(P, k, device=dev) for k in (3, 4, 3)]
out = R.render(R.make_view(sc["cam"], dev), pc, R.PIPE_DEFAULT, torch.zeros(3, device=dev), z[0], z[1], z[2],
               is_smooth_gaussian_features=True, want_viewspace_grad=False)
out["render_gaussian_features"].sum().backward()
torch.cuda.synchronize()
print("feature state ok", float(pc._gaussian_features.grad.abs().sum()))
