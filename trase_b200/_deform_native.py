"""Native (sm_100a) evaluation of DeformNetwork: fused positional encoding, eight
TF32 tensor-core layers with bias+ReLU epilogues, one fused head GEMM, and a
hand-written backward (dgrad with the ReLU derivative and the bias gradient in
the epilogue, wgrad straight from the row-major activations).

Mirrors DeformNetwork.forward (utils/time_utils.py:106-131) for the default
non-blender, non-6DoF model.  Numerics: TF32 products (10-bit mantissa
operands, rounded to nearest by every producer), fp32 accumulation; the
reference runs fp32 SGEMM.  xyz and t carry no gradient (train.py:202-204
detaches xyz).
"""
from __future__ import annotations

import ctypes
import weakref

import torch

from . import _cuda, _lib

W = 256
FUSED_FORWARD = True   # module switch: False runs the layer-by-layer GEMM chain (csrc/mlp_gemm.cu)
FUSED_BACKWARD = True  # same for the data-gradient chain of the backward pass (needs FUSED_FORWARD)
# When every hidden-layer parameter's `.grad` is a view of a flat gradient arena that opted in
# (trase_b200.distributed.GradArena marks its parameters), the weight- and bias-gradient kernels add straight into
# it (they accumulate anyway) and backward() returns None for those inputs: same result as autograd's
# `grad += new` without 16 tiny add kernels and a zero-filled buffer.  This bypasses autograd's own accumulation, so
# it is OPT-IN per parameter (only GradArena parameters) and never taken when a parameter carries gradient hooks or
# when `.grad` was replaced since the arena was built; everything else gets the gradients returned the normal way.
# Restriction that comes with the opt-in: drive an arena-backed network with `.backward()`, not with
# torch.autograd.grad (which must not touch `.grad`).  Module switch: False disables it everywhere.
DIRECT_GRAD_ACCUM = True
# The activations and layer gradients that the two whole-network kernels write to HBM are read by exactly one
# consumer, the weight-gradient products, which are bound by streaming them.  With this switch they are written
# (and read) as BF16 — the layer chain itself keeps its TF32 operands — which halves that traffic; weight
# gradients then carry bf16 rounding noise (8 mantissa bits per factor, averaged over the M rows of the sum;
# measured against fp32 in tests/test_gpu_mlp.py next to the TF32 yardstick).  False: fp32 tensors, TF32 products.
BF16_WGRAD = True
# with bf16 operands: all weight-gradient products of a backward pass in one launch (trase_b200_wgrad_bf16_multi)
MULTI_WGRAD = True


def _direct_ok(p, dev, align=16) -> bool:
    g = p.grad
    return (getattr(p, "_trase_arena_grad_ptr", None) is not None and g is not None and
            g.data_ptr() == p._trase_arena_grad_ptr and g.is_contiguous() and g.dtype == torch.float32 and
            g.device == dev and g.data_ptr() % align == 0 and g.shape == p.shape and
            not getattr(p, "_post_accumulate_grad_hooks", None) and not getattr(p, "_backward_hooks", None))


# ---- optional per-stage timing (bench.py's `stages`): CUDA event pairs on the launching stream around the forward
# kernel, the backward data-gradient chain and the weight-gradient launches; off by default
_prof_on = False
_prof_live = []          # (stage, start_event, end_event)


def profile_enable(on: bool) -> None:
    global _prof_on
    _prof_on = bool(on)


class _Span:
    def __init__(self, stage):
        self.stage = stage

    def __enter__(self):
        if _prof_on:
            self.a = torch.cuda.Event(enable_timing=True)
            self.a.record()
        return self

    def __exit__(self, *exc):
        if _prof_on:
            b = torch.cuda.Event(enable_timing=True)
            b.record()
            _prof_live.append((self.stage, self.a, b))
        return False


def profile_read() -> dict:
    """{stage: (total ms, spans)} for mlp_fwd / mlp_bwd_chain / mlp_wgrad since the last read."""
    out = {}
    for stage, a, b in _prof_live:
        b.synchronize()
        tot, cnt = out.get(stage, (0.0, 0))
        out[stage] = (tot + a.elapsed_time(b), cnt + 1)
    _prof_live.clear()
    return out


def tf32_round(w: torch.Tensor) -> torch.Tensor:
    """Round-to-nearest to TF32 (10 mantissa bits), like cvt.rna.tf32.f32."""
    return ((w.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


def _stream(dev):
    return _cuda.raw_stream(dev)


def _gemm(lib, A, B, D, bias=None, relu=False, mask_in=None, mask_out=None, colsum=None, round_tf32=True):
    M, K = A.shape
    N = B.shape[0]
    _lib.check(lib.trase_b200_gemm_tf32(
        A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), D.data_ptr(), D.stride(0), M, N, K,
        bias.data_ptr() if bias is not None else None, int(relu),
        mask_in.data_ptr() if mask_in is not None else None, mask_out.data_ptr() if mask_out is not None else None,
        colsum.data_ptr() if colsum is not None else None, int(round_tf32), _stream(A.device)))


def _wgrad(lib, P, Q, D):
    """D[i, j] += sum_m P[m, i] Q[m, j]; Q wider than 256 columns goes in chunks."""
    M, I = P.shape
    J = Q.shape[1]
    s = _stream(P.device)
    for j0 in range(0, J, 256):
        j1 = min(J, j0 + 256)
        _lib.check(lib.trase_b200_wgrad_tf32(P.data_ptr(), P.stride(0), I, Q[:, j0:].data_ptr(), Q.stride(0), j1 - j0,
                                             M, D[:, j0:].data_ptr(), D.stride(0), s))


def _wgrad16(lib, P16, Q16, D):
    """D[i, j] += sum_m P16[m, i] Q16[m, j] with bf16 operands (csrc/mlp_gemm.cu wgrad_bf16_kernel)."""
    M, I = P16.shape
    J = D.shape[1]
    _lib.check(lib.trase_b200_wgrad_bf16(P16.data_ptr(), P16.stride(0), I, Q16.data_ptr(), Q16.stride(0), J, M,
                                         D.data_ptr(), D.stride(0), _stream(P16.device)))


def _wgrad16_multi(lib, jobs):
    """All products of `jobs` = [(P16, Q16, D), ...] in one launch (csrc/mlp_gemm.cu wgrad_bf16_multi_kernel)."""
    n = len(jobs)
    if n == 0:
        return
    M = jobs[0][0].shape[0]
    vp, ip = ctypes.c_void_p * n, ctypes.c_int * n
    _lib.check(lib.trase_b200_wgrad_bf16_multi(
        n, vp(*[P.data_ptr() for P, _, _ in jobs]), ip(*[P.stride(0) for P, _, _ in jobs]),
        ip(*[P.shape[1] for P, _, _ in jobs]), vp(*[Q.data_ptr() for _, Q, _ in jobs]),
        ip(*[Q.stride(0) for _, Q, _ in jobs]), ip(*[D.shape[1] for _, _, D in jobs]), M,
        vp(*[D.data_ptr() for _, _, D in jobs]), ip(*[D.stride(0) for _, _, D in jobs]), _stream(jobs[0][0].device)))


def _to_bf16(lib, src, cols, ld_dst):
    """[M][ld_dst] bf16 copy of the first `cols` columns of the fp32 matrix `src` (padding columns zero)."""
    M = src.shape[0]
    dst = torch.empty((M, ld_dst), dtype=torch.bfloat16, device=src.device)
    _lib.check(lib.trase_b200_f32_to_bf16(M, cols, src.data_ptr(), src.stride(0), dst.data_ptr(), ld_dst,
                                          _stream(src.device)))
    return dst


def _round_transpose(lib, Ws, sizes, skip, E, want_rounded, f32, s):
    """Wr[i] = TF32-rounded weight (only if want_rounded), WrT[i] = its transpose for i > 0, restricted to the hidden
    input columns for the layer behind the skip; all in one flat buffer, one small kernel per layer."""
    nl = len(sizes)
    n_r = sum(r * c for r, c in sizes) if want_rounded else 0
    n_t = sum(r * (c - E if i - 1 == skip else c) for i, (r, c) in enumerate(sizes) if i > 0)
    wbuf = torch.empty(n_r + n_t, **f32)
    Wr, WrT, off = [], [None] * nl, 0
    for i, (r, c) in enumerate(sizes):
        wr = None
        if want_rounded:
            wr = wbuf[off:off + r * c].view(r, c)
            off += r * c
        wt = None
        if i > 0:
            c0 = E if i - 1 == skip else 0
            wt = wbuf[off:off + r * (c - c0)].view(c - c0, r)
            off += r * (c - c0)
            WrT[i] = wt
        if wr is not None or wt is not None:
            _lib.check(lib.trase_b200_mlp_round_transpose(
                Ws[i].detach().contiguous().data_ptr(), r, c, E if (i > 0 and i - 1 == skip) else 0,
                (c - E) if (i > 0 and i - 1 == skip) else (c if i > 0 else 0),
                wr.data_ptr() if wr is not None else None, wt.data_ptr() if wt is not None else None, s))
        Wr.append(wr)
    return Wr, WrT


# packed forward operands of frozen networks: (stream, shape, (data_ptr, version) of every parameter) -> tensors
_PACK_CACHE: dict = {}


class _DeformMLP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, t, Lx, Lt, skip, keep, *params):
        lib = _lib.load()
        nl = (len(params) - 6) // 2
        Ws, bs = params[0:2 * nl:2], params[1:2 * nl:2]
        heads = params[2 * nl:]
        dev = x.device
        M = x.shape[0]
        E = 3 + 6 * Lx + 1 + 2 * Lt                       # 84
        with _cuda.on_device(dev):
            s = _stream(dev)
            x = x.detach().contiguous().float()
            t = t.detach().float()
            t_stride = 0 if (t.numel() == 1 or t.stride(0) == 0) else 1
            t_c = t if t_stride == 0 else t.contiguous()
            f32 = dict(dtype=torch.float32, device=dev)
            sizes = [(w.shape[0], w.shape[1]) for w in Ws]
            fused = FUSED_FORWARD and lib.trase_b200_mlp_fused_pack_floats(nl, skip, E) > 0 and \
                all(r == W for r, _ in sizes)
            # TF32-rounded weights (forward operand of the layer chain) and their transposes (backward operand);
            # the fused forward packs its own operand, so there the transposes are left to backward()
            Wr, WrT = (None, [None] * nl) if fused else _round_transpose(lib, Ws, sizes, skip, E, True, f32, s)
            emb, emb16 = None, None
            bf16 = fused and keep and BF16_WGRAD and FUSED_BACKWARD and lib.trase_b200_mlp_fused_pack_bwd_floats(nl) > 0
            if fused:
                # the positional encoding is the longest of the small preparatory kernels: launched first, it
                # covers the host time needed to launch the others
                emb = torch.empty((M, E), **f32)
                if bf16:      # backward reads the embedding only as a weight-gradient operand: a bf16 copy for it
                    emb16 = torch.empty((M, (E + 7) // 8 * 8), dtype=torch.bfloat16, device=dev)
                    _lib.check(lib.trase_b200_deform_embed_bf16(M, x.data_ptr(), t_c.data_ptr(), t_stride, Lx, Lt,
                                                                emb.data_ptr(), emb.stride(0), emb16.data_ptr(),
                                                                emb16.stride(0), s))
                else:
                    _lib.check(lib.trase_b200_deform_embed(M, x.data_ptr(), t_c.data_ptr(), t_stride, Lx, Lt,
                                                           emb.data_ptr(), emb.stride(0), s))
            # An inference call (torch.no_grad: FEATURE state, style transfer, rendering) on weights that have not
            # changed since the last one reuses the packed operands instead of re-packing them per view.  Training
            # calls always pack: their weights change every step.
            pack_key = None
            cached = None
            if fused and not keep and all(isinstance(p, torch.nn.Parameter) for p in params):
                # keyed on the parameter OBJECTS (checked through weak references: an address or id recycled by a new
                # network never matches), their storage and their version counters (any in-place update bumps them)
                pack_key = (s, nl, skip, E) + tuple((id(p), p.data_ptr(), p._version) for p in params)
                cached = _PACK_CACHE.get(pack_key)
                if cached is not None and not all(r() is p for r, p in zip(cached[0], params)):
                    cached = None
                if cached is not None:
                    cached = cached[1:]
            # fused path: the kernel writes the three heads as the contiguous tensors handed on; the layer-by-layer
            # path produces the packed [M][16] rows and slices them
            out16 = None if fused else torch.empty((M, 16), **f32)
            heads_out = (torch.empty((M, 3), **f32), torch.empty((M, 4), **f32), torch.empty((M, 3), **f32)) \
                if fused else None
            if cached is not None:
                hb, wpack, ball = cached
                wh, whT, bh = hb[:16 * W].view(16, W), hb[16 * W:32 * W].view(W, 16), hb[32 * W:]
            else:
                hb = torch.empty(16 * W * 2 + 16, **f32)
                wh, whT, bh = hb[:16 * W].view(16, W), hb[16 * W:32 * W].view(W, 16), hb[32 * W:]
                _lib.check(lib.trase_b200_mlp_pack_heads(*[p.detach().contiguous().data_ptr() for p in heads], W,
                                                         wh.data_ptr(), whT.data_ptr(), bh.data_ptr(), s))
            if fused:
                # one persistent kernel carries each 128-row tile through all layers (csrc/mlp_fused.cu)
                if cached is None:
                    wpack = torch.empty(lib.trase_b200_mlp_fused_pack_floats(nl, skip, E), **f32)
                    wptrs = (ctypes.c_void_p * nl)(*[w.detach().contiguous().data_ptr() for w in Ws])
                    _lib.check(lib.trase_b200_mlp_pack_fused(nl, skip, E, wptrs, wh.data_ptr(), wpack.data_ptr(), s))
                    ball = torch.cat([b.detach().reshape(-1) for b in bs] + [bh])
                    if pack_key is not None:
                        if len(_PACK_CACHE) >= 4:
                            _PACK_CACHE.clear()
                        _PACK_CACHE[pack_key] = (tuple(weakref.ref(p) for p in params), hb, wpack, ball)
                if not keep:
                    # inference (torch.no_grad / frozen network): no activations or masks leave the chip
                    with _Span("mlp_fwd"):
                        _lib.check(lib.trase_b200_mlp_forward_fused(M, nl, skip, E, emb.data_ptr(), emb.stride(0),
                                                                    wpack.data_ptr(), ball.data_ptr(), None, None,
                                                                    None, *[h.data_ptr() for h in heads_out], s))
                    return heads_out
                masks = torch.empty((nl, M, W // 32), dtype=torch.int32, device=dev)
                acts_all = torch.empty((nl, M, W), dtype=torch.bfloat16 if bf16 else torch.float32, device=dev)
                fwd = lib.trase_b200_mlp_forward_fused_bf16 if bf16 else lib.trase_b200_mlp_forward_fused
                with _Span("mlp_fwd"):
                    _lib.check(fwd(M, nl, skip, E, emb.data_ptr(), emb.stride(0), wpack.data_ptr(), ball.data_ptr(),
                                   acts_all.data_ptr(), masks.data_ptr(), None, *[h.data_ptr() for h in heads_out], s))
                acts = list(acts_all.unbind(0))
                skipbuf = emb16 if bf16 else emb
            else:
                masks = torch.empty((nl, M, W // 32), dtype=torch.int32, device=dev)
                # skip buffer [M, E + W]: embedding in [:, :E], output of layer `skip` in [:, E:]
                skipbuf = torch.empty((M, E + W), **f32)
                _lib.check(lib.trase_b200_deform_embed(M, x.data_ptr(), t_c.data_ptr(), t_stride, Lx, Lt,
                                                       skipbuf.data_ptr(), skipbuf.stride(0), s))
                acts = []
                inp = skipbuf[:, :E]
                for i in range(nl):
                    out = skipbuf[:, E:] if i == skip else torch.empty((M, W), **f32)
                    _gemm(lib, inp, Wr[i], out, bias=bs[i].detach(), relu=True, mask_out=masks[i])
                    acts.append(out)
                    inp = skipbuf if i == skip else out
                _gemm(lib, acts[-1], wh, out16, bias=bh, round_tf32=False)
        ctx.save_for_backward(skipbuf, whT, masks, *acts, *(Ws if fused else [w for w in WrT if w is not None]))
        ctx.meta = (nl, skip, E, M, sizes, fused)
        ctx.bf16 = bool(fused and acts[0].dtype == torch.bfloat16)
        # the parameter objects themselves: backward looks their .grad up (DIRECT_GRAD_ACCUM)
        hp = list(params[:2 * nl])
        ctx.hidden_params = hp if all(isinstance(p, torch.nn.Parameter) for p in hp) else None
        hd = list(params[2 * nl:])
        ctx.head_params = hd if all(isinstance(p, torch.nn.Parameter) for p in hd) else None
        if heads_out is not None:
            return heads_out
        return out16[:, 0:3].contiguous(), out16[:, 3:7].contiguous(), out16[:, 7:10].contiguous()

    @staticmethod
    def backward(ctx, g_xyz, g_rot, g_scale):
        lib = _lib.load()
        nl, skip, E, M, sizes, fused = ctx.meta
        saved = ctx.saved_tensors
        skipbuf, whT, masks = saved[0], saved[1], saved[2]
        acts = saved[3:3 + nl]
        dev = skipbuf.device
        f32 = dict(dtype=torch.float32, device=dev)
        with _cuda.on_device(dev):
            s = _stream(dev)
            # the fused chain packs its operand from the raw weights the fused forward kept
            bf16 = ctx.bf16                 # the forward wrote bf16 activations: the fused chain + bf16 wgrad follow
            fused_bwd = (bf16 or FUSED_BACKWARD) and fused and lib.trase_b200_mlp_fused_pack_bwd_floats(nl) > 0
            # bf16 operands + the fused chain: every product of the pass is queued and runs as ONE launch at the end
            multi = [] if (bf16 and MULTI_WGRAD and fused_bwd) else None
            if multi is not None:
                def wgrad(_lib_, P, Q, D):
                    multi.append((P, Q, D))
            else:
                wgrad = _wgrad16 if bf16 else _wgrad
            raw_w = list(saved[3 + nl:]) if fused else None
            if fused_bwd:
                WrT = None
            elif fused:
                WrT = _round_transpose(lib, raw_w, sizes, skip, E, False, f32, s)[1]
            else:
                WrT = [None] + list(saved[3 + nl:])
            # gradient targets: the parameters' own .grad where possible (DIRECT_GRAD_ACCUM), else one zeroed buffer
            hidden = ctx.hidden_params
            direct = DIRECT_GRAD_ACCUM and hidden is not None and all(_direct_ok(p, dev) for p in hidden)
            gbuf = torch.zeros((0 if direct else sum(r * c for r, c in sizes) + nl * W) + W * 16 + 16, **f32)
            off = 0
            if direct:
                grads_w, grads_b = [p.grad for p in hidden[0::2]], [p.grad for p in hidden[1::2]]
            else:
                grads_w, grads_b = [], []
                for r, c in sizes:
                    grads_w.append(gbuf[off:off + r * c].view(r, c))
                    off += r * c
                for _ in range(nl):
                    grads_b.append(gbuf[off:off + W])
                    off += W
            dwhT = gbuf[off:off + W * 16].view(W, 16)
            dbh = gbuf[off + W * 16:off + W * 16 + 16]
            dout = torch.empty((M, 16), **f32)
            gs = [None if g is None else g.contiguous() for g in (g_xyz, g_rot, g_scale)]
            gp = [None if g is None else g.data_ptr() for g in gs]
            # bf16 path: the head gradient travels as bf16 hi + lo parts ([M][32]) and its product lands in d32
            dout16 = torch.empty((M, 32), dtype=torch.bfloat16, device=dev) if bf16 else None
            d32 = torch.zeros((W, 32), **f32) if bf16 else None
            if bf16:
                _lib.check(lib.trase_b200_mlp_pack_dout_bf16(M, *gp, dout.data_ptr(), dout16.data_ptr(), dbh.data_ptr(), s))
            else:
                _lib.check(lib.trase_b200_mlp_pack_dout(M, *gp, dout.data_ptr(), dbh.data_ptr(), s))
            # heads: dWh^T = h_last^T dout
            with _Span("mlp_wgrad"):
                if bf16:
                    wgrad(lib, acts[-1], dout16, d32)
                else:
                    wgrad(lib, acts[-1], dout, dwhT)
            if fused_bwd:
                # the whole data-gradient chain in one persistent kernel (csrc/mlp_fused.cu): dzs[L] = dZ of
                # layer nl-1-L, bias gradients straight into their targets
                wpack_b = torch.empty(lib.trase_b200_mlp_fused_pack_bwd_floats(nl), **f32)
                wptrs = (ctypes.c_void_p * nl)(*[w.detach().contiguous().data_ptr() for w in raw_w])
                _lib.check(lib.trase_b200_mlp_pack_fused_bwd(nl, skip, E, whT.data_ptr(), wptrs, wpack_b.data_ptr(), s))
                dzs = torch.empty((nl, M, W), dtype=torch.bfloat16 if bf16 else torch.float32, device=dev)
                bptrs = (ctypes.c_void_p * nl)(*[g.data_ptr() for g in grads_b])
                bwd = lib.trase_b200_mlp_backward_fused_bf16 if bf16 else lib.trase_b200_mlp_backward_fused
                with _Span("mlp_bwd_chain"):
                    _lib.check(bwd(M, nl, dout.data_ptr(), wpack_b.data_ptr(), masks.data_ptr(), dzs.data_ptr(), bptrs, s))
            else:
                # dZ_{nl-1} = (dout Wh) * relu'(h_last); its column sums are the bias gradient of layer nl-1
                dz = torch.empty((M, W), **f32)
                _gemm(lib, dout, whT, dz, mask_in=masks[nl - 1], colsum=grads_b[nl - 1])
            span = _Span("mlp_wgrad" if fused_bwd else "mlp_bwd_unfused")   # fused chain: this loop is wgrad only
            span.__enter__()
            for i in range(nl - 1, -1, -1):
                if fused_bwd:
                    dz = dzs[nl - 1 - i]
                if fused and i > 0 and i - 1 == skip:
                    # the skip layer's input was never concatenated: embedding columns, then the hidden ones
                    wgrad(lib, dz, skipbuf, grads_w[i][:, :E])
                    wgrad(lib, dz, acts[i - 1], grads_w[i][:, E:])
                else:
                    inp = skipbuf[:, :E] if i == 0 else (skipbuf if i - 1 == skip else acts[i - 1])
                    wgrad(lib, dz, inp, grads_w[i])
                if i > 0 and not fused_bwd:
                    dz_prev = torch.empty((M, W), **f32)
                    _gemm(lib, dz, WrT[i], dz_prev, mask_in=masks[i - 1], colsum=grads_b[i - 1])
                    dz = dz_prev
            if multi is not None:
                _wgrad16_multi(lib, multi)
            span.__exit__()
            out = [None, None, None, None, None, None]
            for i in range(nl):
                out += [None, None] if direct else [grads_w[i], grads_b[i]]
            heads = ctx.head_params
            if DIRECT_GRAD_ACCUM and heads is not None and all(_direct_ok(p, dev, align=4) for p in heads):
                # the six head gradients straight into their arena slots: one launch instead of the hi + lo add and
                # six accumulations by autograd
                src, ld, halves = (d32, 32, 2) if bf16 else (dwhT, 16, 1)
                _lib.check(lib.trase_b200_mlp_heads_grad_accumulate(
                    src.data_ptr(), ld, halves, dbh.data_ptr(), *[p.grad.data_ptr() for p in heads], W, s))
                out += [None] * 6
            else:
                if bf16:
                    torch.add(d32[:, :16], d32[:, 16:], out=dwhT)   # hi-part and lo-part products of the head gradient
                dwh = dwhT.t()
                out += [dwh[0:3], dbh[0:3], dwh[3:7], dbh[3:7], dwh[7:10], dbh[7:10]]
        return tuple(out)


def _blender_params(net, t):
    """is_blender variant (utils/time_utils.py:78-80, 108-109): the time embedding goes through `timenet`
    (13 -> 256 -> 30) before it joins the 63 position columns.  One view has ONE time stamp, so that 30-vector tau is
    the same for every row and its share of the two layers that read the embedding is a constant:
        W[:, 63:93] @ tau   added to that layer's bias.
    The kernels then run the ordinary network on a 64-column embedding (63 position columns + one zero column, zero
    weight column to match); timenet itself and the bias folding are two tiny torch ops whose autograd carries the
    gradients back to timenet and to the weight columns that were folded away."""
    from .deform import positional_encoding
    if not (t.numel() == 1 or t.stride(0) == 0 or t.shape[0] == 1):
        if not bool((t == t[:1]).all()):
            raise NotImplementedError("native is_blender DeformNetwork: one time stamp per call (the reference always "
                                      "passes fid.unsqueeze(0).expand(N, -1), train.py:196, render.py)")
    tau = net.timenet(positional_encoding(t[:1].reshape(1, 1), net.t_multires))[0]          # [30]
    nx = 3 + 6 * net.multires
    nt = tau.shape[0]
    params = []
    for i, lin in enumerate(net.linear):
        Wt, b = lin.weight, lin.bias
        if i == 0 or (i - 1) in net.skips:
            zero_col = Wt.new_zeros(Wt.shape[0], 1)
            b = b + Wt[:, nx:nx + nt] @ tau
            Wt = torch.cat([Wt[:, :nx], zero_col, Wt[:, nx + nt:]], dim=1)
        params += [Wt, b]
    return params


def deform_forward(net, x, t):
    if net.is_6dof:
        raise NotImplementedError("native DeformNetwork path covers the non-6DoF models")
    if not x.is_cuda:
        raise RuntimeError("trase_b200: the native DeformNetwork path runs on CUDA tensors only")
    if net.is_blender:
        params = _blender_params(net, t)
        for head in (net.gaussian_warp, net.gaussian_rotation, net.gaussian_scaling):
            params += [head.weight, head.bias]
        keep = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        # Lt = 0: the embedding kernel writes the 63 position columns and the raw "t" column, which is fed a zero
        return _DeformMLP.apply(x, x.new_zeros(1, 1), net.multires, 0, net.skips[0], keep, *params)
    params = []
    for lin in net.linear:
        params += [lin.weight, lin.bias]
    for head in (net.gaussian_warp, net.gaussian_rotation, net.gaussian_scaling):
        params += [head.weight, head.bias]
    # grad mode is always off inside Function.forward, so decide here whether anything must be kept for backward
    keep = torch.is_grad_enabled() and any(p.requires_grad for p in params)
    return _DeformMLP.apply(x, t, net.multires, net.t_multires, net.skips[0], keep, *params)
