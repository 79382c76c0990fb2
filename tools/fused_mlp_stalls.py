"""Stall-cycle attribution of the two whole-network MLP kernels in TRAINING mode (bf16 side outputs on):
cycles per CTA the MMA-issuing thread and one epilogue warp spent waiting on each barrier.
The counters are compiled out of the shipped library (they cost 4-5 % of the kernels); build the instrumented one first:
    TRASE_B200_EXTRA_NVCC_FLAGS=-DTB_FU_PROF=1 TRASE_B200_LIB_OUT=libtrase_b200_prof.so python -m trase_b200.build"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trase_b200 import _lib, _deform_native as dn
from trase_b200.deform import DeformNetwork

torch.manual_seed(0)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 300000
net = DeformNetwork(backend="native").cuda()
x = (torch.rand(M, 3, device="cuda") - 0.5) * 3
t = torch.full((M, 1), 0.37, device="cuda")
_prof = os.path.join(os.path.dirname(_lib.LIB_PATH), "libtrase_b200_prof.so")
if os.path.exists(_prof):
    _lib.LIB_PATH = _prof
else:
    print("note: %s not built; the stall counters of the shipped library read 0" % _prof)
lib = _lib.load()
lib.trase_b200_mlp_fused_set_prof.argtypes = [ctypes.c_void_p]
buf = torch.zeros(148 * 8, dtype=torch.int64, device="cuda")


def show(tag):
    for name, sl in (("leader", slice(0, 148, 2)), ("follower", slice(1, 148, 2))):
        b = buf.view(148, 8)[sl].double().mean(0).tolist()
        print(tag, name, "cycles/CTA: total %.0f | MMA thread waits: tmem_empty %.0f a_ready %.0f w_full %.0f store %.0f | "
              "epilogue warp 4 waits: tmem_full %.0f h_free %.0f | MMA issue + commits %.0f" % tuple(b[:8]))


MODES = [bool(int(m)) for m in sys.argv[2].split(',')] if len(sys.argv) > 2 else [True, False]
for mode in MODES:
    dn.BF16_WGRAD = mode
    for _ in range(2):
        o = net(x, t)
        torch.autograd.backward(list(o), [torch.ones_like(a) for a in o])
    torch.cuda.synchronize()
    lib.trase_b200_mlp_fused_set_prof(buf.data_ptr())
    buf.zero_()
    o = net(x, t)
    torch.cuda.synchronize()
    show("bf16=%d forward " % mode)
    buf.zero_()
    torch.autograd.backward(list(o), [torch.ones_like(a) for a in o])
    torch.cuda.synchronize()
    show("bf16=%d backward" % mode)
    lib.trase_b200_mlp_fused_set_prof(None)
    dn.profile_enable(True)
    for _ in range(5):
        o = net(x, t)
        torch.autograd.backward(list(o), [torch.ones_like(a) for a in o])
    torch.cuda.synchronize()
    print("bf16=%d" % mode, {k: round(v[0] / v[1], 4) if k != "mlp_wgrad" else round(v[0] / 5, 4) for k, v in dn.profile_read().items()})
    dn.profile_enable(False)
