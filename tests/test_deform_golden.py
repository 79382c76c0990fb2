"""DeformNetwork mirror against golden vectors produced by the reference's own
module (tests/golden/make_golden_mlp.py).  CPU: the torch backend must match to
fp32 rounding; GPU: the native TF32 path within TF32 accuracy."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR
from trase_b200.deform import DeformNetwork


def _load(tag, device="cpu", backend="torch"):
    g = np.load(os.path.join(GOLDEN_DIR, "deform_mlp.npz"))
    torch.manual_seed(int(g[tag + "_seed"]))
    net = DeformNetwork(is_blender=(tag == "blender"), backend=backend)
    # same construction order as the reference => same seeded parameters, under the same names
    for n, p in net.named_parameters():
        assert abs(p.detach().double().sum().item() - float(g[f"{tag}_psum_{n}"])) <= 1e-9, n
    net = net.to(device)
    x, t = torch.from_numpy(g[tag + "_x"]).to(device), torch.from_numpy(g[tag + "_t"]).to(device)
    go = [torch.from_numpy(g[f"{tag}_go{i}"]).to(device) for i in range(3)]
    return g, net, x, t, go


@pytest.mark.parametrize("tag", ["default", "blender"])
def test_torch_backend_matches_reference_module(tag):
    g, net, x, t, go = _load(tag)
    outs = net(x, t)
    torch.autograd.backward(list(outs), go)
    for i, o in enumerate(outs):
        assert np.abs(o.detach().numpy() - g[f"{tag}_out{i}"]).max() <= 1e-6
    for n, p in net.named_parameters():
        if f"{tag}_grad_{n}" in g.files:
            ref = g[f"{tag}_grad_{n}"]
            assert np.abs(p.grad.numpy() - ref).max() <= 1e-5 * max(1.0, np.abs(ref).max()), n


def test_unsupported_variants_fail_loudly():
    with pytest.raises(NotImplementedError):
        DeformNetwork(is_6dof=True)
    net = DeformNetwork(backend="native")
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        net(torch.zeros(4, 3), torch.zeros(4, 1))


@pytest.mark.gpu
def test_native_backend_matches_reference_within_tf32(native_lib):
    dev = torch.device("cuda:0")
    g, net, x, t, go = _load("default", dev, backend="native")
    outs = net(x, t)
    torch.autograd.backward(list(outs), go)
    for i, o in enumerate(outs):
        ref = g[f"default_out{i}"]
        assert np.abs(o.detach().cpu().numpy() - ref).max() <= 5e-3 * max(1.0, np.abs(ref).max())
    for n, p in net.named_parameters():
        if f"default_grad_{n}" not in g.files:
            continue
        ref = g[f"default_grad_{n}"]
        err = np.linalg.norm(p.grad.cpu().numpy() - ref) / max(np.linalg.norm(ref), 1e-20)
        # 10-bit-mantissa products, fp32 accumulation: on this random-init net torch's own TF32 (cuBLAS)
        # path sits at 2-4 % from fp32 for the early layers (tests/test_gpu_mlp.py measures that yardstick)
        assert err <= 6e-2, (n, err)
