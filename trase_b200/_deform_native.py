"""Native (sm_100a) evaluation of DeformNetwork: fused positional encoding, eight
TF32 tensor-core layers with bias+ReLU epilogues, one fused head GEMM, and a
hand-written backward (dgrad with the ReLU derivative and the bias gradient in
the epilogue, wgrad straight from the row-major activations).

Mirrors DeformNetwork.forward (utils/time_utils.py:106-131) for the default
non-blender, non-6DoF model.  Numerics: TF32 products (10-bit mantissa
operands), fp32 accumulation; the reference runs fp32 SGEMM.  xyz and t carry
no gradient (train.py:202-204 detaches xyz).
"""
from __future__ import annotations

import torch

from . import _lib

W = 256


def tf32_round(w: torch.Tensor) -> torch.Tensor:
    """Round-to-nearest to TF32 (10 mantissa bits), like cvt.rna.tf32.f32."""
    return ((w.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


def _gemm(lib, A, B, D, bias=None, relu=False, mask_in=None, mask_out=None, colsum=None, round_tf32=True):
    M, K = A.shape
    N = B.shape[0]
    _lib.check(lib.trase_b200_gemm_tf32(
        A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), D.data_ptr(), D.stride(0), M, N, K,
        bias.data_ptr() if bias is not None else None, int(relu),
        mask_in.data_ptr() if mask_in is not None else None, mask_out.data_ptr() if mask_out is not None else None,
        colsum.data_ptr() if colsum is not None else None, int(round_tf32),
        torch.cuda.current_stream(A.device).cuda_stream))


def _wgrad(lib, P, Q, D):
    """D[i, j] += sum_m P[m, i] Q[m, j]; Q wider than 256 columns goes in chunks."""
    M, I = P.shape
    J = Q.shape[1]
    s = torch.cuda.current_stream(P.device).cuda_stream
    for j0 in range(0, J, 256):
        j1 = min(J, j0 + 256)
        _lib.check(lib.trase_b200_wgrad_tf32(P.data_ptr(), P.stride(0), I, Q[:, j0:].data_ptr(), Q.stride(0), j1 - j0,
                                             M, D[:, j0:].data_ptr(), D.stride(0), s))


class _DeformMLP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, t, Lx, Lt, skip, *params):
        lib = _lib.load()
        nl = (len(params) - 6) // 2
        Ws, bs = params[0:2 * nl:2], params[1:2 * nl:2]
        heads_w, heads_b = params[2 * nl::2], params[2 * nl + 1::2]
        dev = x.device
        M = x.shape[0]
        E = 3 + 6 * Lx + 1 + 2 * Lt                       # 84
        with torch.cuda.device(dev):
            x = x.detach().contiguous().float()
            t = t.detach().float()
            t_stride = 0 if (t.numel() == 1 or t.stride(0) == 0) else 1
            t_c = t if t_stride == 0 else t.contiguous()
            f32 = dict(dtype=torch.float32, device=dev)
            # skip buffer [M, E + W]: embedding in [:, :E], output of layer `skip` in [:, E:]
            skipbuf = torch.empty((M, E + W), **f32)
            _lib.check(lib.trase_b200_deform_embed(M, x.data_ptr(), t_c.data_ptr(), t_stride, Lx, Lt, skipbuf.data_ptr(),
                                                   skipbuf.stride(0), torch.cuda.current_stream(dev).cuda_stream))
            emb = skipbuf[:, :E]
            Wr = [tf32_round(w.detach()) for w in Ws]
            acts, masks = [], []
            inp = emb
            for i in range(nl):
                out = skipbuf[:, E:] if i == skip else torch.empty((M, W), **f32)
                mk = torch.empty((M, W // 32), dtype=torch.int32, device=dev)
                _gemm(lib, inp, Wr[i], out, bias=bs[i].detach(), relu=True, mask_out=mk)
                acts.append(out)
                masks.append(mk)
                inp = skipbuf if i == skip else out
            # the three heads as one [16, W] GEMM (10 real rows)
            wh = torch.zeros((16, W), **f32)
            bh = torch.zeros((16,), **f32)
            wh[0:3], wh[3:7], wh[7:10] = heads_w[0].detach(), heads_w[1].detach(), heads_w[2].detach()
            bh[0:3], bh[3:7], bh[7:10] = heads_b[0].detach(), heads_b[1].detach(), heads_b[2].detach()
            out16 = torch.empty((M, 16), **f32)
            wh = tf32_round(wh)
            _gemm(lib, acts[-1], wh, out16, bias=bh, round_tf32=False)
        ctx.save_for_backward(skipbuf, wh, *acts, *Wr, *masks)
        ctx.meta = (nl, skip, E, M)
        return out16[:, 0:3].contiguous(), out16[:, 3:7].contiguous(), out16[:, 7:10].contiguous()

    @staticmethod
    def backward(ctx, g_xyz, g_rot, g_scale):
        lib = _lib.load()
        nl, skip, E, M = ctx.meta
        saved = ctx.saved_tensors
        skipbuf, wh = saved[0], saved[1]
        acts = saved[2:2 + nl]
        Ws = saved[2 + nl:2 + 2 * nl]
        masks = saved[2 + 2 * nl:2 + 3 * nl]
        dev = skipbuf.device
        f32 = dict(dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            dout = torch.zeros((M, 16), **f32)
            if g_xyz is not None:
                dout[:, 0:3] = g_xyz
            if g_rot is not None:
                dout[:, 3:7] = g_rot
            if g_scale is not None:
                dout[:, 7:10] = g_scale
            dbh = dout.sum(0)
            dout = tf32_round(dout)
            # heads: dWh^T = h_last^T dout ; db = colsum(dout)
            dwhT = torch.zeros((W, 16), **f32)
            _wgrad(lib, acts[-1], dout, dwhT)
            grads_w = [None] * nl
            grads_b = [None] * nl
            # dZ_{nl-1} = (dout Wh) * relu'(h_last)
            dz = torch.empty((M, W), **f32)
            db = torch.zeros((W,), **f32)
            _gemm(lib, dout, wh.t().contiguous(), dz, mask_in=masks[-1], colsum=db)
            for i in range(nl - 1, -1, -1):
                grads_b[i] = db
                inp = skipbuf[:, :E] if i == 0 else (skipbuf if i - 1 == skip else acts[i - 1])
                dw = torch.zeros((W, inp.shape[1]), **f32)
                _wgrad(lib, dz, inp, dw)
                grads_w[i] = dw
                if i > 0:
                    wt = Ws[i][:, E:] if i - 1 == skip else Ws[i]      # only the hidden part carries gradient
                    dz_prev = torch.empty((M, W), **f32)
                    db = torch.zeros((W,), **f32)
                    _gemm(lib, dz, wt.t().contiguous(), dz_prev, mask_in=masks[i - 1], colsum=db)
                    dz = dz_prev
            dwh = dwhT.t()
            out = [None, None, None, None, None]
            for i in range(nl):
                out += [grads_w[i], grads_b[i]]
            out += [dwh[0:3].contiguous(), dbh[0:3].contiguous(), dwh[3:7].contiguous(), dbh[3:7].contiguous(),
                    dwh[7:10].contiguous(), dbh[7:10].contiguous()]
        return tuple(out)


def deform_forward(net, x, t):
    if net.is_blender or net.is_6dof:
        raise NotImplementedError("native DeformNetwork path covers the default (non-blender, non-6DoF) model")
    if not x.is_cuda:
        raise RuntimeError("trase_b200: the native DeformNetwork path runs on CUDA tensors only")
    params = []
    for lin in net.linear:
        params += [lin.weight, lin.bias]
    for head in (net.gaussian_warp, net.gaussian_rotation, net.gaussian_scaling):
        params += [head.weight, head.bias]
    return _DeformMLP.apply(x, t, net.multires, net.t_multires, net.skips[0], *params)
