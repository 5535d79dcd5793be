"""Thin torch-tensor wrappers over the C ABI plus the autograd glue of the hot path.

torch is used for device memory, streams and autograd bookkeeping only; every FLOP goes through
libds2_b200.so.  Tensors handed to the C ABI must be CUDA + contiguous."""
from __future__ import annotations

import math
import weakref
from ctypes import c_void_p
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import DS2_BF16, DS2_F32, check, lib

RELU_CAP = 20.0                                   # src/custom_ops.py:71
SBN_SCALE = 1.0 / math.sqrt(1.0 + 1e-5)           # src/custom_ops.py:184-199 as written


def _p(t: Optional[torch.Tensor]):
    return None if t is None else c_void_p(t.data_ptr())


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return DS2_F32
    if t.dtype == torch.bfloat16:
        return DS2_BF16
    raise ValueError("ds2: unsupported dtype %s (float32 / bfloat16 only)" % t.dtype)


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("deepspeech_b200 ops need CUDA tensors (no CPU fallback)")


_ws_cache = {}


def _workspace(nbytes: int, device, tag: str) -> torch.Tensor:
    key = (tag, device)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


# ------------------------------------------------------------------------------------ GEMM
def gemm_nt(A: torch.Tensor, B: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False, M=None, N=None, K=None,
            lda=None, ldb=None, out: Optional[torch.Tensor] = None, out_dtype=torch.float32, alpha: float = 1.0,
            beta: float = 0.0, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """C[M,N] = alpha * A . B^T + bias + beta*C.  A is [M,K] (a_mn False) or [K,M] (a_mn True); B is
    [N,K] or [K,N].  A/B may be 2-D row slices of larger buffers (lda/ldb = row pitch)."""
    _need_cuda(A, B)
    if M is None:
        M = A.shape[1] if a_mn else A.shape[0]
    if K is None:
        K = A.shape[0] if a_mn else A.shape[1]
    if N is None:
        N = B.shape[1] if b_mn else B.shape[0]
    lda = A.stride(0) if lda is None else lda
    ldb = B.stride(0) if ldb is None else ldb
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=A.device)
    ldc = out.stride(0)
    use_tc = (A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and lda % 8 == 0 and ldb % 8 == 0
              and A.data_ptr() % 16 == 0 and B.data_ptr() % 16 == 0 and K >= 8)
    if use_tc:
        check(lib.ds2_gemm_bf16(_p(A), lda, int(a_mn), _p(B), ldb, int(b_mn), _p(out), _dt(out), ldc, M, N, K,
                                alpha, beta, _p(bias), _stream()), "ds2_gemm_bf16")
    else:
        sam, sak = (1, lda) if a_mn else (lda, 1)
        sbn, sbk = (1, ldb) if b_mn else (ldb, 1)
        check(lib.ds2_gemm_generic(_p(A), _dt(A), sam, sak, _p(B), _dt(B), sbn, sbk, _p(out), _dt(out), ldc, M, N, K,
                                   alpha, beta, _p(bias), _stream()), "ds2_gemm_generic")
    return out


def gemm_wgrad(A, B, M, N, K, lda, ldb, out, alpha=1.0):
    """out[M,N] (fp32) = alpha * A^T.B with both operands MN-major ([K,M] / [K,N] row-major): the weight-gradient
    shape (few output tiles, very long K).  Tensor cores with automatic split-K when the operands allow it."""
    if (A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and lda % 8 == 0 and ldb % 8 == 0
            and A.data_ptr() % 16 == 0 and B.data_ptr() % 16 == 0 and K >= 8 and out.dtype == torch.float32):
        check(lib.ds2_gemm_bf16_splitk(_p(A), lda, 1, _p(B), ldb, 1, _p(out), out.stride(0), M, N, K, alpha, None, 0, _stream()),
              "ds2_gemm_bf16_splitk")
        return out
    return gemm_nt(A, B, a_mn=True, b_mn=True, M=M, N=N, K=K, lda=lda, ldb=ldb, out=out, alpha=alpha)


def cast(x: torch.Tensor, dtype) -> torch.Tensor:
    if x.dtype == dtype:
        return x
    out = torch.empty(x.shape, dtype=dtype, device=x.device)
    check(lib.ds2_cast(_p(x.contiguous()), _dt(x), _p(out), _dt(out), x.numel(), _stream()), "ds2_cast")
    return out


# bf16 copies of the fp32 master weights (and U^T for the backward recurrence) are made once per weight VERSION, not once
# per step: keyed by storage address + shape, valid while the tensor's autograd version and the global epoch (bumped by
# optimizers that update parameters through the C ABI, which autograd cannot see) are unchanged.
_wcache = {}
_weight_epoch = 0


def invalidate_weight_cache():
    global _weight_epoch
    _weight_epoch += 1
    _wcache.clear()


def _cached(w: torch.Tensor, kind: str, make):
    # only long-lived parameters are cached, and an entry belongs to one tensor OBJECT (weak reference): a temporary
    # that happens to reuse a freed tensor's address must never hit
    if not isinstance(w, torch.nn.Parameter):
        return make()
    key = (id(w), kind)
    tag = (w._version, _weight_epoch, w.data_ptr())
    e = _wcache.get(key)
    if e is not None and e[0]() is w and e[1] == tag:
        return e[2]
    out = make()
    _wcache[key] = (weakref.ref(w), tag, out)
    return out


def cached_bf16(w: torch.Tensor) -> torch.Tensor:
    if w.dtype == torch.bfloat16:
        return w
    return _cached(w, "bf16", lambda: cast(w.contiguous(), torch.bfloat16))


def cached_ut_bf16(U: torch.Tensor) -> torch.Tensor:
    """U [ND,H,H] -> bf16 [ND,H,H] with ut[d] = U[d]^T (the resident operand of the backward recurrence)."""
    def make():
        ND, H, _ = U.shape
        Uc = U.contiguous()
        ut = torch.empty((ND, H, H), dtype=torch.bfloat16, device=U.device)
        for d in range(ND):
            check(lib.ds2_transpose(_p(Uc[d]), _dt(Uc), H, H, H, _p(ut[d]), DS2_BF16, H, _stream()), "ds2_transpose")
        return ut
    return _cached(U, "ut", make)


def add2(a: torch.Tensor, b: torch.Tensor, out_dtype) -> torch.Tensor:
    out = torch.empty(a.shape, dtype=out_dtype, device=a.device)
    check(lib.ds2_add2(_p(a), _p(b), _dt(a), _p(out), _dt(out), a.numel(), _stream()), "ds2_add2")
    return out


def colsum(x2d: torch.Tensor, R: int, N: int, ld: int) -> torch.Tensor:
    out = torch.empty(N, dtype=torch.float32, device=x2d.device)
    check(lib.ds2_colsum(_p(x2d), _dt(x2d), R, N, ld, _p(out), 0, _stream()), "ds2_colsum")
    return out


# ------------------------------------------------------------------------------------ relux
class _ReluxFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cap):
        xc = x.contiguous()
        y = torch.empty_like(xc)
        has_cap = cap is not None
        check(lib.ds2_relux_fwd(_p(xc), _p(y), xc.numel(), float(cap or 0.0), int(has_cap), _stream()), "ds2_relux_fwd")
        ctx.save_for_backward(xc)
        ctx.cap = cap
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dyc = dy.contiguous()
        dx = torch.empty_like(x)
        check(lib.ds2_relux_bwd(_p(x), _p(dyc), _p(dx), x.numel(), float(ctx.cap or 0.0), int(ctx.cap is not None), _stream()),
              "ds2_relux_bwd")
        return dx, None


def relux(x: torch.Tensor, capping=None) -> torch.Tensor:
    _need_cuda(x)
    if x.dtype != torch.float32:
        raise ValueError("relux: float32 tensors only")
    return _ReluxFn.apply(x, capping)


# ------------------------------------------------------------------------------------ BN (+clip), channel-last
def _bn_scratch(device):
    return _workspace(2 * 256 * 8, device, "bn").view(torch.float64)


def bn_stats(x: torch.Tensor, C: int, eps: float) -> Tuple[torch.Tensor, torch.Tensor]:
    R = x.numel() // C
    mean = torch.empty(C, dtype=torch.float32, device=x.device)
    rstd = torch.empty(C, dtype=torch.float32, device=x.device)
    check(lib.ds2_bn_stats(_p(x), R, C, eps, _p(_bn_scratch(x.device)), _p(mean), _p(rstd), _stream()), "ds2_bn_stats")
    return mean, rstd


def bn_stats_from_sums(sums: torch.Tensor, count: int, C: int, eps: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """mean / rstd from the per-channel sums a producer kernel accumulated in its epilogue (f64 [2C]: sum, sum of squares)."""
    mean = torch.empty(C, dtype=torch.float32, device=sums.device)
    rstd = torch.empty(C, dtype=torch.float32, device=sums.device)
    check(lib.ds2_bn_finalize(_p(sums), float(count), C, eps, _p(mean), _p(rstd), _stream()), "ds2_bn_finalize")
    return mean, rstd


class _BnClipFn(torch.autograd.Function):
    """y = clip(gamma * (x - mean_batch) * rstd_batch + beta) over a channel-last [B,T,F,C] tensor;
    the output is written with element strides `omap` = (sb, st, sf, sc) into a tensor of shape
    `oshape` / dtype `odtype` (NHWC itself, or the RNN layout)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps, cap, oshape, omap, odtype, sums=None):
        B, T, Fq, C = x.shape
        mean, rstd = bn_stats(x, C, eps) if sums is None else bn_stats_from_sums(sums, B * T * Fq, C, eps)
        out = torch.empty(oshape, dtype=odtype, device=x.device)
        clip = cap is not None
        check(lib.ds2_bn_clip_fwd(_p(x), B, T, Fq, C, _p(mean), _p(rstd), _p(gamma), _p(beta), float(cap or 0.0), int(clip),
                                  _p(out), _dt(out), *omap, _stream()), "ds2_bn_clip_fwd")
        ctx.save_for_backward(x, gamma, beta, mean, rstd)
        ctx.cap, ctx.omap = cap, omap
        return out

    @staticmethod
    def backward(ctx, dout):
        x, gamma, beta, mean, rstd = ctx.saved_tensors
        B, T, Fq, C = x.shape
        dout = dout.contiguous()
        dx = torch.empty_like(x)
        dgamma = torch.empty_like(gamma)
        dbeta = torch.empty_like(beta)
        check(lib.ds2_bn_clip_bwd(_p(x), B, T, Fq, C, _p(mean), _p(rstd), _p(gamma), _p(beta), float(ctx.cap or 0.0),
                                  int(ctx.cap is not None), _p(dout), _dt(dout), *ctx.omap, _p(_bn_scratch(x.device)),
                                  _p(dx), _p(dgamma), _p(dbeta), _stream()), "ds2_bn_clip_bwd")
        return dx, dgamma, dbeta, None, None, None, None, None, None


def bn_clip_nhwc(x, gamma, beta, eps, cap=None):
    """channel-last in, channel-last out (fp32)."""
    B, T, Fq, C = x.shape
    omap = (T * Fq * C, Fq * C, C, 1)
    return _BnClipFn.apply(x.contiguous(), gamma, beta, eps, cap, tuple(x.shape), omap, torch.float32, None)


def bn_clip_to_rnn_layout(x, gamma, beta, eps, cap, feature_order: str, odtype, sums=None):
    """channel-last [B,T,F,C] in, time-major [T,B,C*F] out; feature = c*F+f ('cf', NCHW model,
    src/deepSpeech_NCHW.py:166-168) or f*C+c ('fc', NHWC model, src/deepSpeech.py:162-164)."""
    B, T, Fq, C = x.shape
    feat = C * Fq
    omap = (feat, B * feat, 1, Fq) if feature_order == "cf" else (feat, B * feat, C, 1)
    return _BnClipFn.apply(x.contiguous(), gamma, beta, eps, cap, (T, B, feat), omap, odtype, sums)


# ------------------------------------------------------------------------------------ conv2d (channel-last, HWIO)
class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, st, sf):
        B, Ti, Fi, Ci = x.shape
        KT, KF, Ci2, Co = w.shape
        if Ci2 != Ci:
            raise ValueError("conv2d: filter in-channels %d != input channels %d" % (Ci2, Ci))
        To, Fo = (Ti - KT) // st + 1, (Fi - KF) // sf + 1
        if To <= 0 or Fo <= 0:
            raise ValueError("conv2d: input %dx%d smaller than the %dx%d filter" % (Ti, Fi, KT, KF))
        y = torch.empty((B, To, Fo, Co), dtype=torch.float32, device=x.device)
        check(lib.ds2_conv2d_fwd(_p(x), _dt(x), _p(w), _p(bias), _p(y), B, Ti, Fi, Ci, Co, KT, KF, st, sf, _stream()),
              "ds2_conv2d_fwd")
        ctx.save_for_backward(x, w)
        ctx.geom = (B, Ti, Fi, Ci, Co, KT, KF, st, sf)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        g = ctx.geom
        dy = dy.contiguous()
        dw = torch.empty_like(w)
        db = torch.empty(g[4], dtype=torch.float32, device=x.device) if ctx.has_bias else None
        check(lib.ds2_conv2d_wgrad(_p(x), _dt(x), _p(dy), _p(dw), _p(db), *g, _stream()), "ds2_conv2d_wgrad")
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(x.shape, dtype=torch.float32, device=x.device)
            check(lib.ds2_conv2d_dgrad(_p(dy), _p(w), _p(dx), *g, _stream()), "ds2_conv2d_dgrad")
            if x.dtype != torch.float32:
                dx = cast(dx, x.dtype)
        return dx, dw, db, None, None


def conv2d_nhwc(x, w, bias, st, sf):
    _need_cuda(x, w)
    return _Conv2dFn.apply(x.contiguous(), w.contiguous(), bias, st, sf)


class _ConvGemmFn(torch.autograd.Function):
    """y = conv(f(x), w) + bias on the tensor cores, f = identity or clip(BN_batch(x)) of the previous
    layer applied while the patch matrix is gathered (ds2_im2col -> ds2_gemm_bf16).  x fp32 channel-last
    [B,Ti,Fi,Ci]; w HWIO fp32; y fp32 [B,To,Fo,Co].  Backward: dW = P^T dy (split-K GEMM), db = colsum,
    dx through dP = dy W^T (GEMM) -> ds2_col2im -> BN/clip backward."""

    @staticmethod
    def forward(ctx, x, w, bias, st, sf, gamma, beta, eps, cap):
        B, Ti, Fi, Ci = x.shape
        KT, KF, _, Co = w.shape
        To, Fo = (Ti - KT) // st + 1, (Fi - KF) // sf + 1
        if To <= 0 or Fo <= 0:
            raise ValueError("conv2d: input %dx%d smaller than the %dx%d filter" % (Ti, Fi, KT, KF))
        K = KT * KF * Ci
        Kp = (K + 7) // 8 * 8
        M = B * To * Fo
        dev = x.device
        has_bn = gamma is not None
        mean = rstd = None
        if has_bn:
            mean, rstd = bn_stats(x, Ci, eps)
        P = torch.empty((M, Kp), dtype=torch.bfloat16, device=dev)
        check(lib.ds2_im2col(_p(x), _p(mean), _p(rstd), _p(gamma), _p(beta), float(cap or 0.0), int(cap is not None), _p(P), Kp,
                             B, Ti, Fi, Ci, KT, KF, st, sf, _stream()), "ds2_im2col")
        wb = torch.zeros((Kp, Co), dtype=torch.bfloat16, device=dev) if Kp != K else torch.empty((Kp, Co), dtype=torch.bfloat16, device=dev)
        check(lib.ds2_cast(_p(w), DS2_F32, _p(wb), DS2_BF16, K * Co, _stream()), "ds2_cast")
        y = gemm_nt(P, wb, b_mn=True, M=M, N=Co, K=Kp, bias=bias)                       # [M,Co] fp32
        ctx.save_for_backward(x, gamma, beta, mean, rstd, P, wb)
        ctx.geom = (B, Ti, Fi, Ci, KT, KF, st, sf)
        ctx.misc = (Co, K, Kp, M, eps, cap, has_bn, bias is not None, tuple(w.shape))
        return y.view(B, To, Fo, Co)

    @staticmethod
    def backward(ctx, dy):
        x, gamma, beta, mean, rstd, P, wb = ctx.saved_tensors
        B, Ti, Fi, Ci, KT, KF, st, sf = ctx.geom
        Co, K, Kp, M, eps, cap, has_bn, has_bias, wshape = ctx.misc
        dev = x.device
        dy2 = dy.contiguous().view(M, Co)
        dyb = cast(dy2, torch.bfloat16)
        dwp = torch.empty((Kp, Co), dtype=torch.float32, device=dev)
        check(lib.ds2_gemm_bf16_splitk(_p(P), Kp, 1, _p(dyb), Co, 1, _p(dwp), Co, Kp, Co, M, 1.0, None, 0, _stream()),
              "ds2_gemm_bf16_splitk")
        dw = dwp[:K].view(wshape)
        db = colsum(dy2, M, Co, Co) if has_bias else None
        dx = dgamma = dbeta = None
        if ctx.needs_input_grad[0]:
            dP = gemm_nt(dyb, wb, M=M, N=Kp, K=Co, out_dtype=torch.bfloat16)            # [M,Kp] = dy . W^T
            da = torch.empty(x.shape, dtype=torch.float32, device=dev)
            check(lib.ds2_col2im(_p(dP), Kp, _p(da), B, Ti, Fi, Ci, KT, KF, st, sf, _stream()), "ds2_col2im")
            del dP
            if has_bn:
                dx = torch.empty_like(x)
                dgamma = torch.empty_like(gamma)
                dbeta = torch.empty_like(beta)
                omap = (Ti * Fi * Ci, Fi * Ci, Ci, 1)
                check(lib.ds2_bn_clip_bwd(_p(x), B, Ti, Fi, Ci, _p(mean), _p(rstd), _p(gamma), _p(beta), float(cap or 0.0),
                                          int(cap is not None), _p(da), DS2_F32, *omap, _p(_bn_scratch(dev)), _p(dx), _p(dgamma),
                                          _p(dbeta), _stream()), "ds2_bn_clip_bwd")
            else:
                dx = da
        return dx, dw, db, None, None, dgamma, dbeta, None, None


def _slack_bf16(shape, device):
    """bf16 tensor of `shape` whose storage extends 64 zeroed bytes past the last element (the window view of
    ds2_conv_window_* reads them with its last box)."""
    n = int(np.prod(shape))
    flat = torch.empty(n + 32, dtype=torch.bfloat16, device=device)
    flat[n:].zero_()
    return flat[:n].view(shape)


def conv_window_padded_dy(dy, B, Ti, Fi, KT, KF, st):
    """dL/dy fp32 [B,To,Fo,32] -> the zero-bordered bf16 copy ds2_conv_window_dgrad / _wgrad read (layout: include/ds2_b200.h)."""
    import ctypes
    ptop, rdy = ctypes.c_int(0), ctypes.c_int(0)
    check(lib.ds2_conv_window_dy_layout(Ti, Fi, KT, KF, st, ctypes.byref(ptop), ctypes.byref(rdy)), "ds2_conv_window_dy_layout")
    dyp = _slack_bf16((B, rdy.value, Fi - KF + 1 + 2 * (KF - 1), 32), dy.device)
    check(lib.ds2_conv_window_pad_cast(_p(dy), _p(dyp), B, Ti, Fi, KT, KF, st, _stream()), "ds2_conv_window_pad_cast")
    return dyp


class _ConvWindowFn(torch.autograd.Function):
    """Same contract as _ConvGemmFn, without a patch matrix: a = f(x) is written once in bf16 and the tensor cores read
    overlapping filter windows of it through a strided TMA view (ds2_conv_window_*; 32 channels in and out, sf = 1)."""

    @staticmethod
    def forward(ctx, x, w, bias, st, gamma, beta, eps, cap, sums_in, want_sums):
        B, Ti, Fi, Ci = x.shape
        KT, KF, _, Co = w.shape
        To, Fo = (Ti - KT) // st + 1, Fi - KF + 1
        if To <= 0 or Fo <= 0:
            raise ValueError("conv2d: input %dx%d smaller than the %dx%d filter" % (Ti, Fi, KT, KF))
        dev = x.device
        has_bn = gamma is not None
        mean = rstd = None
        a = _slack_bf16((B, Ti, Fi, Ci), dev)
        if has_bn:
            mean, rstd = bn_stats(x, Ci, eps) if sums_in is None else bn_stats_from_sums(sums_in, B * Ti * Fi, Ci, eps)
            omap = (Ti * Fi * Ci, Fi * Ci, Ci, 1)
            check(lib.ds2_bn_clip_fwd(_p(x), B, Ti, Fi, Ci, _p(mean), _p(rstd), _p(gamma), _p(beta), float(cap or 0.0),
                                      int(cap is not None), _p(a), DS2_BF16, *omap, _stream()), "ds2_bn_clip_fwd")
        else:
            check(lib.ds2_cast(_p(x), DS2_F32, _p(a), DS2_BF16, x.numel(), _stream()), "ds2_cast")
        K = KT * KF * Ci
        wt = torch.empty((Co, K), dtype=torch.bfloat16, device=dev)
        wd = torch.empty((Ci, K), dtype=torch.bfloat16, device=dev)
        check(lib.ds2_conv_window_weights(_p(w), _p(wt), _p(wd), KT, KF, _stream()), "ds2_conv_window_weights")
        y = torch.empty((B, To, Fo, Co), dtype=torch.float32, device=dev)
        sums = torch.empty(2 * Co, dtype=torch.float64, device=dev) if want_sums else None
        check(lib.ds2_conv_window_fwd(_p(a), _p(wt), _p(bias), _p(y), _p(sums), B, Ti, Fi, KT, KF, st, _stream()), "ds2_conv_window_fwd")
        ctx.save_for_backward(x, gamma, beta, mean, rstd, a, wd)
        ctx.geom = (B, Ti, Fi, Ci, KT, KF, st, To, Fo, Co)
        ctx.misc = (eps, cap, has_bn, bias is not None, tuple(w.shape))
        if want_sums:
            ctx.mark_non_differentiable(sums)
            return y, sums
        return y

    @staticmethod
    def backward(ctx, dy, *_unused):
        x, gamma, beta, mean, rstd, a, wd = ctx.saved_tensors
        B, Ti, Fi, Ci, KT, KF, st, To, Fo, Co = ctx.geom
        eps, cap, has_bn, has_bias, wshape = ctx.misc
        dev = x.device
        dy = cast(dy.contiguous(), torch.float32)
        dyp = conv_window_padded_dy(dy, B, Ti, Fi, KT, KF, st)
        dw = torch.empty(wshape, dtype=torch.float32, device=dev)
        check(lib.ds2_conv_window_wgrad(_p(a), _p(dyp), _p(dw), B, Ti, Fi, KT, KF, st, _stream()), "ds2_conv_window_wgrad")
        db = colsum(dy.view(B * To * Fo, Co), B * To * Fo, Co, Co) if has_bias else None
        dx = dgamma = dbeta = None
        if ctx.needs_input_grad[0]:
            da = torch.empty(x.shape, dtype=torch.float32, device=dev)
            check(lib.ds2_conv_window_dgrad(_p(dyp), _p(wd), _p(da), B, Ti, Fi, KT, KF, st, _stream()), "ds2_conv_window_dgrad")
            if has_bn:
                dx = torch.empty_like(x)
                dgamma = torch.empty_like(gamma)
                dbeta = torch.empty_like(beta)
                omap = (Ti * Fi * Ci, Fi * Ci, Ci, 1)
                check(lib.ds2_bn_clip_bwd(_p(x), B, Ti, Fi, Ci, _p(mean), _p(rstd), _p(gamma), _p(beta), float(cap or 0.0),
                                          int(cap is not None), _p(da), DS2_F32, *omap, _p(_bn_scratch(dev)), _p(dx), _p(dgamma),
                                          _p(dbeta), _stream()), "ds2_bn_clip_bwd")
            else:
                dx = da
        return dx, dw, db, None, dgamma, dbeta, None, None, None, None


class _Conv1FusedFn(torch.autograd.Function):
    """y = conv(x, w) + bias for a one-channel input and 32 filters, the patch rows built in shared memory
    (ds2_conv1_fused_*); also returns the per-channel sums of y and y^2 its epilogue accumulated (for the batch norm that
    follows).  x fp32 [B,Ti,Fi,1]; no gradient flows into x."""

    @staticmethod
    def forward(ctx, x, w, bias, st, sf):
        B, Ti, Fi, _ = x.shape
        KT, KF, _, Co = w.shape
        To, Fo = (Ti - KT) // st + 1, (Fi - KF) // sf + 1
        y = torch.empty((B, To, Fo, Co), dtype=torch.float32, device=x.device)
        sums = torch.empty(2 * Co, dtype=torch.float64, device=x.device)
        check(lib.ds2_conv1_fused_fwd(_p(x), _p(w), _p(bias), _p(y), _p(sums), B, Ti, Fi, KT, KF, st, sf, _stream()), "ds2_conv1_fused_fwd")
        ctx.save_for_backward(x)
        ctx.geom = (B, Ti, Fi, KT, KF, st, sf)
        ctx.misc = (tuple(w.shape), bias is not None)
        ctx.mark_non_differentiable(sums)
        return y, sums

    @staticmethod
    def backward(ctx, dy, _dsums):
        (x,) = ctx.saved_tensors
        wshape, has_bias = ctx.misc
        dy = cast(dy.contiguous(), torch.float32)
        dw = torch.empty(wshape, dtype=torch.float32, device=x.device)
        db = torch.empty(wshape[3], dtype=torch.float32, device=x.device) if has_bias else None
        check(lib.ds2_conv1_fused_wgrad(_p(x), _p(dy), _p(dw), _p(db), *ctx.geom, _stream()), "ds2_conv1_fused_wgrad")
        return None, dw, db, None, None


USE_CONV_WINDOW = True
USE_CONV1_FUSED = True


def conv1_fused_takes(x, w, st, sf) -> bool:
    KT, KF, Ci, Co = w.shape
    return bool(USE_CONV1_FUSED and x.dtype == torch.float32 and x.is_cuda and
                lib.ds2_conv1_fused_supported(Ci, Co, x.shape[1], x.shape[2], KT, KF, st, sf))


def conv1_fused(x, w, bias, st, sf):
    """-> (y fp32 [B,To,Fo,32], sums f64 [64]) ; x fp32 [B,Ti,Fi,1]."""
    _need_cuda(x, w)
    return _Conv1FusedFn.apply(x.contiguous(), w.contiguous(), bias, st, sf)


def conv2d_gemm(x, w, bias, st, sf, gamma=None, beta=None, eps=1e-3, cap=None, sums_in=None, want_sums=False):
    """Tensor-core convolution of f(x): f = clip(BN(x)) when gamma/beta are given, identity otherwise.
    32-channel layers with sf = 1 take the implicit (window) form, everything else the explicit patch matrix.
    sums_in: per-channel sums of x from its producer's epilogue (instead of a statistics pass over x); want_sums: also
    return the sums of the output (window form only: -> (y, sums); the patch-matrix form returns (y, None))."""
    _need_cuda(x, w)
    if x.dtype != torch.float32:
        raise ValueError("conv2d_gemm: fp32 input expected")
    KT, KF, Ci, Co = w.shape
    if USE_CONV_WINDOW and lib.ds2_conv_window_supported(Ci, Co, x.shape[2], KT, KF, sf):
        return _ConvWindowFn.apply(x.contiguous(), w.contiguous(), bias, st, gamma, beta, eps, cap, sums_in, want_sums)
    y = _ConvGemmFn.apply(x.contiguous(), w.contiguous(), bias, st, sf, gamma, beta, eps, cap)
    return (y, None) if want_sums else y


# ------------------------------------------------------------------------------------ recurrent stack
def rnn_path_dtype(B: int, H: int, ND: int, want_bf16: bool):
    """bf16 (persistent tcgen05 kernel) when requested and the shape is taken, else fp32."""
    if want_bf16 and lib.ds2_rnn_relu_persistent_supported(B, H, ND):
        return torch.bfloat16
    return torch.float32


import os as _os0
USE_FP32_LAYER_SUM = _os0.environ.get("DS2_NO_FP32_LAYER_SUM") is None and _os0.environ.get("DS2_RNN_V1") is None


def _relu_recurrence_fwd(wx, Ug, bias, lens, hs, T, B, H, ND, reverse_mask, ws, wsb, scale=SBN_SCALE, cap=RELU_CAP):
    """Runs the forward recurrence into hs [ND,T,B,H] and returns the tensor the layer output fw + bw
    (src/custom_ops.py:95) is formed from: on the bf16 path the kernel's fp32 copy of the hidden states (values before
    their rounding to bf16: the layer output is then rounded once instead of three times), else hs itself."""
    if hs.dtype == torch.bfloat16 and USE_FP32_LAYER_SUM:
        hs32 = torch.empty((ND, T, B, H), dtype=torch.float32, device=hs.device)
        check(lib.ds2_rnn_relu_fwd_ex(_p(wx), scale, _p(Ug), _p(bias), _p(lens), _p(hs), _p(hs32), T, B, H, ND,
                                      reverse_mask, cap, _p(ws), wsb, _stream()), "ds2_rnn_relu_fwd_ex")
        return hs32
    check(lib.ds2_rnn_relu_fwd(_p(wx), scale, _p(Ug), _p(bias), _p(lens), _p(hs), _dt(hs), T, B, H, ND,
                               reverse_mask, cap, _p(ws), wsb, _stream()), "ds2_rnn_relu_fwd")
    return hs


class _BiRNNLayerFn(torch.autograd.Function):
    """One (bi)directional clipped-ReLU layer over the whole sequence:
         wx = x . Wcat^T            (one batched GEMM, all timesteps, both directions)
         hs = recurrence(wx, U, b)  (persistent kernel or per-step kernels)
         out = sum_d hs[d]          (src/custom_ops.py:95)
       x [T,B,I]; Wcat [ND*H, I]; U [ND,H,H]; bias [ND,H]; lens int32 [B] (device).
       Gradients flow in fp32 between layers (x.grad has x's dtype)."""

    @staticmethod
    def forward(ctx, x, Wcat, U, bias, lens, reverse_mask, want_bf16, out_dtype, scale=SBN_SCALE, cap=RELU_CAP):
        T, B, I = x.shape
        ND, H = U.shape[0], U.shape[1]
        dev = x.device
        cdt = rnn_path_dtype(B, H, ND, want_bf16)
        if cdt == torch.bfloat16:
            xg, Wg, Ug = cast(x.contiguous(), torch.bfloat16), cached_bf16(Wcat), cached_bf16(U)
        else:
            xg, Wg, Ug = cast(x.contiguous(), torch.float32), Wcat, U
        wx = gemm_nt(xg.view(T * B, I), Wg)                                       # [TB, ND*H] fp32
        hs = torch.empty((ND, T, B, H), dtype=cdt, device=dev)
        wsb = lib.ds2_rnn_workspace_bytes(T, B, H, ND, _dt(hs))
        ws = _workspace(wsb, dev, "rnn") if wsb else None
        src = _relu_recurrence_fwd(wx, Ug, bias, lens, hs, T, B, H, ND, reverse_mask, ws, wsb, scale, cap)
        if ND == 2:
            out = add2(src[0], src[1], out_dtype)
        else:
            out = cast(src[0], out_dtype)
        ctx.save_for_backward(xg, Wg, Ug, hs, lens)
        ctx.meta = (T, B, I, H, ND, reverse_mask, x.dtype)
        ctx.scale_cap = (float(scale), float(cap))
        return out

    @staticmethod
    def backward(ctx, dout):
        xg, Wg, Ug, hs, lens = ctx.saved_tensors
        T, B, I, H, ND, reverse_mask, xdtype = ctx.meta
        scale, cap = ctx.scale_cap
        dev = xg.device
        cdt = hs.dtype
        dout = cast(dout.contiguous(), torch.float32)
        dz = torch.empty((T * B, ND * H), dtype=cdt, device=dev)
        wsb = lib.ds2_rnn_workspace_bytes(T, B, H, ND, _dt(hs))
        ws = _workspace(wsb, dev, "rnn") if wsb else None
        db = torch.empty((ND, H), dtype=torch.float32, device=dev)
        check(lib.ds2_rnn_relu_bwd(_p(dout), _p(Ug), _p(lens), _p(hs), _p(dz), _p(db), _dt(hs), T, B, H, ND, reverse_mask,
                                   cap, _p(ws), wsb, _stream()), "ds2_rnn_relu_bwd")
        # dW = scale * dz^T x ; dx = scale * dz W   (both directions in one GEMM each)
        dW = torch.empty((ND * H, I), dtype=torch.float32, device=dev)
        gemm_wgrad(dz, xg.view(T * B, I), ND * H, I, T * B, ND * H, I, dW, alpha=scale)    # [ND*H, I]
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm_nt(dz, Wg, b_mn=True, alpha=scale).view(T, B, I)                   # [TB, I] fp32
            if xdtype != torch.float32:
                dx = cast(dx, xdtype)
        # dU[d] = sum_t dz[d,t]^T h[d,t_prev]
        dU = torch.empty((ND, H, H), dtype=torch.float32, device=dev)
        for d in range(ND):
            rev = (reverse_mask >> d) & 1
            if T > 1:
                a_off = (0 if rev else B * ND * H) + d * H           # dz rows t = 1.. (fw) / 0..T-2 (bw)
                A = dz.view(-1)[a_off:]
                Bm = hs[d].reshape(T * B, H)[(B if rev else 0):]
                gemm_wgrad(A, Bm, H, H, (T - 1) * B, ND * H, H, dU[d])
            else:
                dU[d].zero_()
        return dx, dW, dU, db, None, None, None, None, None, None


def birnn_layer(x, Wcat, U, bias, lens, reverse_mask: int, want_bf16: bool, out_dtype, scale: float = SBN_SCALE,
                cap: float = RELU_CAP):
    """scale multiplies W.x (the reference's seq-BN constant by default), cap clips the ReLU (3e38: plain ReLU, the
    MkldnnRNNRelu cell of src/mkldnn_rnn_op.py)."""
    _need_cuda(x, Wcat, U, bias, lens)
    return _BiRNNLayerFn.apply(x, Wcat.contiguous(), U.contiguous(), bias.contiguous(), lens, reverse_mask, want_bf16, out_dtype,
                               float(scale), float(cap))


def gru_path_dtype(B: int, H: int, ND: int, want_bf16: bool):
    """bf16 (persistent tcgen05 GRU kernel) when requested and the shape is taken, else fp32."""
    if want_bf16 and lib.ds2_gru_persistent_supported(B, H, ND):
        return torch.bfloat16
    return torch.float32


class _BiGRULayerFn(torch.autograd.Function):
    """GRU analogue of _BiRNNLayerFn (build-defined, see include/ds2_b200.h).
       x [T,B,I]; Wcat [ND*3H, I]; U [ND,3H,H]; bw, bu [ND,3H].  bf16 path: tcgen05 GEMMs for W x, dW, dx, dU and the
       persistent recurrence kernels; fp32 path: strided GEMMs and per-step kernels."""

    @staticmethod
    def forward(ctx, x, Wcat, U, bw, bu, lens, reverse_mask, want_bf16, out_dtype):
        T, B, I = x.shape
        ND, H = U.shape[0], U.shape[2]
        dev = x.device
        cdt = gru_path_dtype(B, H, ND, want_bf16)
        hs = torch.empty((ND, T, B, H), dtype=cdt, device=dev)
        gates = torch.empty((ND, T, B, 4 * H), dtype=cdt, device=dev)
        if cdt == torch.bfloat16:
            xg, Wg, Ug = cast(x.contiguous(), torch.bfloat16), cached_bf16(Wcat), cached_bf16(U)
            wx = gemm_nt(xg.view(T * B, I), Wg)                                    # [TB, ND*3H] fp32
            wsb = lib.ds2_gru_workspace_bytes(T, B, H, ND)
            ws = _workspace(wsb, dev, "gru")
            check(lib.ds2_gru_persistent_fwd(_p(wx), SBN_SCALE, _p(Ug), _p(bw), _p(bu), _p(lens), _p(hs), _p(gates), T, B, H,
                                             ND, reverse_mask, _p(ws), wsb, _stream()), "ds2_gru_persistent_fwd")
        else:
            xg, Wg, Ug = cast(x.contiguous(), torch.float32), Wcat, U
            wx = gemm_nt(xg.view(T * B, I), Wg)
            check(lib.ds2_gru_fwd(_p(wx), SBN_SCALE, _p(Ug), _p(bw), _p(bu), _p(lens), _p(hs), _p(gates), DS2_F32, T, B, H, ND,
                                  reverse_mask, _stream()), "ds2_gru_fwd")
        out = add2(hs[0], hs[1], out_dtype) if ND == 2 else cast(hs[0], out_dtype)
        ctx.save_for_backward(xg, Wg, Ug, hs, gates, lens)
        ctx.meta = (T, B, I, H, ND, reverse_mask, x.dtype)
        return out

    @staticmethod
    def backward(ctx, dout):
        xg, Wg, Ug, hs, gates, lens = ctx.saved_tensors
        T, B, I, H, ND, reverse_mask, xdtype = ctx.meta
        dev = xg.device
        cdt = hs.dtype
        G = ND * 3 * H
        dout = cast(dout.contiguous(), torch.float32)
        dgi = torch.empty((T * B, G), dtype=cdt, device=dev)
        dgh = torch.empty((T * B, G), dtype=cdt, device=dev)
        if cdt == torch.bfloat16:
            wsb = lib.ds2_gru_workspace_bytes(T, B, H, ND)
            ws = _workspace(wsb, dev, "gru")
            dbw = torch.empty((ND, 3 * H), dtype=torch.float32, device=dev)
            dbu = torch.empty((ND, 3 * H), dtype=torch.float32, device=dev)
            check(lib.ds2_gru_persistent_bwd(_p(dout), _p(Ug), _p(lens), _p(hs), _p(gates), _p(dgi), _p(dgh), _p(dbw), _p(dbu), T, B,
                                             H, ND, reverse_mask, _p(ws), wsb, _stream()), "ds2_gru_persistent_bwd")
        else:
            dhz = torch.zeros((ND, B, H), dtype=torch.float32, device=dev)
            check(lib.ds2_gru_bwd(_p(dout), _p(Ug), _p(lens), _p(hs), _p(gates), _p(dgi), _p(dgh), _p(dhz), DS2_F32, T, B, H,
                                  ND, reverse_mask, _stream()), "ds2_gru_bwd")
        if cdt != torch.bfloat16:
            dbw = colsum(dgi, T * B, G, G).view(ND, 3 * H)
            dbu = colsum(dgh, T * B, G, G).view(ND, 3 * H)
        dW = torch.empty((G, I), dtype=torch.float32, device=dev)
        gemm_wgrad(dgi, xg.view(T * B, I), G, I, T * B, G, I, dW, alpha=SBN_SCALE)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm_nt(dgi, Wg, b_mn=True, alpha=SBN_SCALE).view(T, B, I)
            if xdtype != torch.float32:
                dx = cast(dx, xdtype)
        # dU[d] = sum_t dgh[d,t]^T h[d,t_prev]
        dU = torch.empty((ND, 3 * H, H), dtype=torch.float32, device=dev)
        for d in range(ND):
            rev = (reverse_mask >> d) & 1
            if T > 1:
                a_off = (0 if rev else B * G) + d * 3 * H
                A = dgh.view(-1)[a_off:]
                Bm = hs[d].reshape(T * B, H)[(B if rev else 0):]
                gemm_wgrad(A, Bm, 3 * H, H, (T - 1) * B, G, H, dU[d])
            else:
                dU[d].zero_()
        return dx, dW, dU, dbw, dbu, None, None, None, None


def bigru_layer(x, Wcat, U, bw, bu, lens, reverse_mask: int, want_bf16: bool, out_dtype):
    _need_cuda(x, Wcat, U, bw, bu, lens)
    return _BiGRULayerFn.apply(x, Wcat.contiguous(), U.contiguous(), bw.contiguous(), bu.contiguous(), lens, reverse_mask,
                               want_bf16, out_dtype)


# ------------------------------------------------------------------------------------ whole recurrent stack
_aux_streams = {}
OVERLAP_WGRAD = True          # weight-gradient GEMMs on a side stream next to the following layer's recurrence
import os as _os
FUSED_BIAS_GRAD = _os.environ.get("DS2_NO_FUSED_DB") is None   # bias gradients accumulated inside the backward recurrence
STACK_GRAD_HOOK = None        # callable(layer, [gradient tensors], [their parameters]) run as soon as a layer's weight gradients are enqueued
                              # (data-parallel reducers start their all-reduce here, see dp.GradAllReducer)
STACK_TRACE = None            # set to a list to collect (tag, layer, start_event, end_event) of the stack backward


def _streams(device):
    """(rec, side): a high-priority stream for the persistent recurrence kernels and a lowest-priority one for the
    work that is off the critical path of the backward pass."""
    key = torch.device(device).index
    st = _aux_streams.get(key)
    if st is None:
        lo, hi = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
        st = (torch.cuda.Stream(device=device, priority=hi), torch.cuda.Stream(device=device, priority=lo))
        _aux_streams[key] = st
    return st


def rec_stack_supported(cell: str, B: int, H: int, ND: int) -> bool:
    if cell == "gru":
        return bool(lib.ds2_gru_persistent_supported(B, H, ND))
    return bool(lib.ds2_rnn_relu_persistent_supported(B, H, ND))


def _rec_ctas(cell: str, B: int, H: int, ND: int) -> int:
    """CTAs (= SMs) a persistent recurrence launch occupies: 4-CTA clusters x 128-unit tiles x directions x chunks."""
    chunks = max(1, B // 32) if cell == "gru" else 1
    ks = 8 if (cell == "gru" and H > 1024) else 4          # the GRU's 8-CTA K split for wide layers
    return ks * ((H + 127) // 128) * ND * chunks


class _RecStackFn(torch.autograd.Function):
    """The 'stacked' wiring (src/custom_ops.py:76-95 applied layer after layer) as ONE autograd node, bf16 path:
         per layer  wx = x W^T (tcgen05 GEMM) -> persistent recurrence -> out = sum_d hs[d]  (bf16, next layer's x)
       The backward is scheduled over three streams: the recurrence of layer l on a high-priority stream, dx of
       layer l (the only GEMM the next recurrence waits for) on the caller's stream, and db/dW/dU of layer l on a
       low-priority stream limited to the SMs the recurrence of layer l-1 leaves free, so they run in its shadow.
       All streams are joined before the gradients are returned.  Gradients between layers stay fp32.
       params: per layer (Wcat [ND*G*H, I], U [ND,G*H,H], bias [ND,G*H]) + (bu [ND,3H]) for the GRU."""

    @staticmethod
    def forward(ctx, x, lens, cell, reverse_mask, out_dtype, L, *params):
        T, B, _ = x.shape
        dev = x.device
        gru = cell == "gru"
        npl = 4 if gru else 3
        saved, metas = [], []
        cur = cast(x.contiguous(), torch.bfloat16)
        for l in range(L):
            W, U = params[l * npl], params[l * npl + 1]
            ND, H = U.shape[0], U.shape[2]
            I = cur.shape[2]
            Wg, Ug = cached_bf16(W), cached_bf16(U)
            wx = gemm_nt(cur.view(T * B, I), Wg)
            hs = torch.empty((ND, T, B, H), dtype=torch.bfloat16, device=dev)
            if gru:
                bw, bu = params[l * npl + 2], params[l * npl + 3]
                gates = torch.empty((ND, T, B, 4 * H), dtype=torch.bfloat16, device=dev)
                wsb = lib.ds2_gru_workspace_bytes(T, B, H, ND)
                ws = _workspace(wsb, dev, "gru")
                check(lib.ds2_gru_persistent_fwd(_p(wx), SBN_SCALE, _p(Ug), _p(bw), _p(bu), _p(lens), _p(hs), _p(gates), T, B,
                                                 H, ND, reverse_mask, _p(ws), wsb, _stream()), "ds2_gru_persistent_fwd")
                saved += [cur, Wg, Ug, hs, gates]
            else:
                bias = params[l * npl + 2]
                wsb = lib.ds2_rnn_workspace_bytes(T, B, H, ND, DS2_BF16)
                ws = _workspace(wsb, dev, "rnn")
                src = _relu_recurrence_fwd(wx, Ug, bias, lens, hs, T, B, H, ND, reverse_mask, ws, wsb)
                saved += [cur, Wg, Ug, hs]
            metas.append((I, H, ND))
            last = l == L - 1
            odt = out_dtype if last else torch.bfloat16
            if gru:
                src = hs
            cur = add2(src[0], src[1], odt) if ND == 2 else cast(src[0], odt)
        ctx.save_for_backward(lens, *saved)
        ctx.meta = (T, B, gru, reverse_mask, L, metas, x.dtype)
        ctx.param_objs = params                     # identities only: handed to STACK_GRAD_HOOK with their gradients
        return cur

    @staticmethod
    def backward(ctx, dout):
        lens, *saved = ctx.saved_tensors
        T, B, gru, reverse_mask, L, metas, xdtype = ctx.meta
        dev = dout.device
        per = 5 if gru else 4
        npl = 4 if gru else 3
        main = torch.cuda.current_stream()
        rec, side = _streams(dev)
        sm_total = torch.cuda.get_device_properties(dev).multi_processor_count
        # Measured (tools/stack_overlap.py, profiles/r1_stack_overlap.txt): GEMMs running next to a recurrence raise its
        # L2 latency; with H=1760 (112 of 148 SMs busy) the recurrence slows by as much as the GEMMs hide, with the
        # GRU at H=800 (56 SMs busy) the overlap gains ~9% of the backward.  Overlap only when most SMs are free.
        rec_sms = max(_rec_ctas("gru" if gru else "relu", B, m[1], m[2]) for m in metas)
        overlap = OVERLAP_WGRAD and L > 1 and (sm_total - rec_sms) * 2 >= sm_total
        if not overlap:
            rec = side = main
        grads = [None] * (L * npl)
        keep = []                                   # everything the side streams touch stays alive until the join
        d_cur = cast(dout.contiguous(), torch.float32)
        side_done = {}                              # layer -> event: its db/dW/dU have finished
        for l in range(L - 1, -1, -1):
            xg, Wg, Ug, hs = saved[l * per:l * per + 4]
            gates = saved[l * per + 4] if gru else None
            I, H, ND = metas[l]
            G = ND * (3 if gru else 1) * H
            # ---- recurrence of layer l (high priority).  It starts together with the side work of layer l+1 and, having
            #      priority, places its clusters first; the side work of layer l+2 must be gone by then, or its CTAs,
            #      scattered over the SMs, could keep some clusters from being placed
            dz = torch.empty((T * B, G), dtype=torch.bfloat16, device=dev)
            dgh = torch.empty((T * B, G), dtype=torch.bfloat16, device=dev) if gru else dz
            # dW, dU, db (, dbu) of a layer live in ONE flat buffer: a data-parallel reducer sends it with one call
            GH = G // ND
            n_w, n_u, n_b = G * I, ND * GH * H, ND * GH
            flat = torch.empty(n_w + n_u + n_b * (2 if gru else 1), dtype=torch.float32, device=dev)
            db = flat[n_w + n_u:n_w + n_u + n_b].view(ND, GH)                           # bias gradients come out of the kernel
            dbu = flat[n_w + n_u + n_b:].view(ND, GH) if gru else None
            keep += [dz, dgh, d_cur, flat]
            if overlap:
                rec.wait_stream(main)
                if l + 2 in side_done:
                    rec.wait_event(side_done[l + 2])
            with torch.cuda.stream(rec):
                if STACK_TRACE is not None:
                    ev0 = torch.cuda.Event(enable_timing=True)
                    ev0.record()
                if gru:
                    wsb = lib.ds2_gru_workspace_bytes(T, B, H, ND)
                    ws = _workspace(wsb, dev, "gru")
                    check(lib.ds2_gru_persistent_bwd(_p(d_cur), _p(Ug), _p(lens), _p(hs), _p(gates), _p(dz), _p(dgh), _p(db if FUSED_BIAS_GRAD else None),
                                                     _p(dbu if FUSED_BIAS_GRAD else None), T,
                                                     B, H, ND, reverse_mask, _p(ws), wsb, _stream()), "ds2_gru_persistent_bwd")
                else:
                    wsb = lib.ds2_rnn_workspace_bytes(T, B, H, ND, DS2_BF16)
                    ws = _workspace(wsb, dev, "rnn")
                    Ut = cached_ut_bf16(ctx.param_objs[l * npl + 1])          # once per weight version, not per step
                    check(lib.ds2_rnn_relu_bwd_ex(_p(d_cur), _p(Ug), _p(Ut), _p(lens), _p(hs), _p(dz), _p(db if FUSED_BIAS_GRAD else None), T, B, H,
                                                  ND, reverse_mask, RELU_CAP, _p(ws), wsb, _stream()), "ds2_rnn_relu_bwd_ex")
                if STACK_TRACE is not None:
                    ev1 = torch.cuda.Event(enable_timing=True)
                    ev1.record()
                    STACK_TRACE.append(("rec", l, ev0, ev1))
            if overlap:
                main.wait_stream(rec)
            # ---- dx of layer l: the critical path
            if l > 0 or ctx.needs_input_grad[0]:
                d_cur = gemm_nt(dz, Wg, b_mn=True, alpha=SBN_SCALE)                      # [TB, I] fp32
            dx_done = torch.cuda.Event()
            dx_done.record(main)
            # ---- db, dW, dU of layer l: off the critical path
            if overlap:
                side.wait_event(dx_done)
                # the SMs the recurrence of layer l-1 leaves free; the last side work has the device to itself
                free = sm_total - _rec_ctas("gru" if gru else "relu", B, metas[l - 1][1], metas[l - 1][2]) if l > 0 else 0
                prev_limit = lib.ds2_set_sm_limit(max(8, free) if l > 0 else 0)
            with torch.cuda.stream(side):
                if STACK_TRACE is not None:
                    ev2 = torch.cuda.Event(enable_timing=True)
                    ev2.record()
                dW = flat[:n_w].view(G, I)
                dU = flat[n_w:n_w + n_u].view(ND, GH, H)
                gemm_wgrad(dz, xg.view(T * B, I), G, I, T * B, G, I, dW, alpha=SBN_SCALE)
                for d in range(ND):
                    rev = (reverse_mask >> d) & 1
                    if T > 1:
                        a_off = (0 if rev else B * G) + d * GH
                        gemm_wgrad(dgh.view(-1)[a_off:], hs[d].reshape(T * B, H)[(B if rev else 0):], GH, H, (T - 1) * B, G, H,
                                   dU[d])
                    else:
                        dU[d].zero_()
                if not FUSED_BIAS_GRAD:
                    db.copy_(colsum(dz, T * B, G, G).view(ND, GH))
                    if gru:
                        dbu.copy_(colsum(dgh, T * B, G, G).view(ND, GH))
                outs = [dW, dU, db]
                if gru:
                    outs.append(dbu)
                taken = False
                if STACK_GRAD_HOOK is not None:
                    # on the stream that produced them: the reducer orders itself after it.  True: the hook has installed
                    # them as the parameters' .grad itself, autograd gets None for them
                    taken = bool(STACK_GRAD_HOOK(l, outs, ctx.param_objs[l * npl:(l + 1) * npl], flat))
                if STACK_TRACE is not None:
                    ev3 = torch.cuda.Event(enable_timing=True)
                    ev3.record()
                    STACK_TRACE.append(("side", l, ev2, ev3))
            if overlap:
                lib.ds2_set_sm_limit(prev_limit)
                side_done[l] = torch.cuda.Event()
                side_done[l].record(side)
                for t in outs:
                    t.record_stream(main)
            grads[l * npl:l * npl + npl] = [None] * len(outs) if taken else outs
        if overlap:
            main.wait_stream(side)
            main.wait_stream(rec)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = d_cur.view(T, B, metas[0][0])
            if xdtype != torch.float32:
                dx = cast(dx, xdtype)
        del keep
        return (dx, None, None, None, None, None, *grads)


def rec_stack(x, lens, cell: str, reverse_mask: int, out_dtype, layer_params):
    """layer_params: list over layers of (Wcat, U, bias) or (Wcat, U, bw, bu) for cell == 'gru'."""
    flat = []
    for lp in layer_params:
        flat += [t.contiguous() for t in lp]
    _need_cuda(x, lens, *flat)
    return _RecStackFn.apply(x, lens, cell, reverse_mask, out_dtype, len(layer_params), *flat)


# ------------------------------------------------------------------------------------ linear
class _LinearFn(torch.autograd.Function):
    """y[R,N] = x[R,K] . W[N,K]^T + b  (fp32 out).  src/deepSpeech_NCHW.py:195-198, custom_ops._linear."""

    @staticmethod
    def forward(ctx, x2d, W, b, want_bf16):
        if want_bf16 and x2d.dtype == torch.bfloat16:
            Wg = cached_bf16(W)
        else:
            Wg = W
            x2d = cast(x2d, torch.float32)
        y = gemm_nt(x2d, Wg, bias=b)
        ctx.save_for_backward(x2d, W)
        ctx.has_bias = b is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x2d, W = ctx.saved_tensors
        dy = dy.contiguous()
        R, K = x2d.shape
        N = W.shape[0]
        dx = None
        tc = x2d.dtype == torch.bfloat16 and R % 8 == 0 and K % 8 == 0
        if tc:
            # tensor cores: dy^T [N,R] in bf16 (one transpose + cast) serves both products
            dyT = torch.empty((N, R), dtype=torch.bfloat16, device=dy.device)
            check(lib.ds2_transpose(_p(dy), DS2_F32, R, N, N, _p(dyT), DS2_BF16, R, _stream()), "ds2_transpose")
        if ctx.needs_input_grad[0]:
            if tc:
                # dx[R,K] = dy . W: dy^T is the MN-major A operand, W [N,K] the MN-major B operand, contraction over N classes
                dx = gemm_nt(dyT, cached_bf16(W), a_mn=True, b_mn=True, M=R, N=K, K=N, out_dtype=torch.bfloat16)
            else:
                dx = gemm_nt(dy, W, b_mn=True)                  # [R,K] = dy[R,N] . W[N,K]
                if x2d.dtype != torch.float32:
                    dx = cast(dx, x2d.dtype)
        if tc:
            # dW[N,K] = dy^T x: dy^T K-major, x the MN-major operand, split-K over the R rows
            dW = torch.empty((N, K), dtype=torch.float32, device=dy.device)
            check(lib.ds2_gemm_bf16_splitk(_p(dyT), R, 0, _p(x2d), K, 1, _p(dW), K, N, K, R, 1.0, None, 0, _stream()),
                  "ds2_gemm_bf16_splitk")
        else:
            dW = gemm_nt(dy, x2d, a_mn=True, b_mn=True)         # [N,K] = dy^T x
        db = colsum(dy, R, N, N) if ctx.has_bias else None
        return dx, dW, db, None


def linear(x2d, W, b, want_bf16=False):
    _need_cuda(x2d, W)
    return _LinearFn.apply(x2d.contiguous(), W.contiguous(), b, want_bf16)


# ------------------------------------------------------------------------------------ CTC
def ctc_prepare_labels(indices, values, dense_shape, num_classes: int, preprocess_collapse_repeated: bool):
    """Host-side label pre-processing of tf.nn.ctc_loss on the SparseTensor built at
    src/deepSpeech_train.py:437: per batch row, in index order, collapse adjacent repeats when asked,
    stop at the first label >= num_classes-1 (a valid label after it is an error).  Returns
    (labels [B,Lmax] int32, label_lens [B] int32, repeats [B] int32) as numpy arrays."""
    idx = np.asarray(indices, dtype=np.int64).reshape(-1, 2)
    vals = np.asarray(values, dtype=np.int64).reshape(-1)
    Bn = int(dense_shape[0])
    if idx.shape[0] != vals.shape[0]:
        raise ValueError("ctc_loss: labels.indices and labels.values disagree in length")
    if idx.shape[0] and (idx[:, 0].min() < 0 or idx[:, 0].max() >= Bn):
        raise ValueError("ctc_loss: label batch index out of range")
    n = idx.shape[0]
    if n == 0:
        return np.zeros((Bn, 1), np.int32), np.zeros(Bn, np.int32), np.zeros(Bn, np.int32)
    # whole-array form of the per-entry scan (no Python loop inside the training step)
    order = np.lexsort((idx[:, 1], idx[:, 0]))
    bs, vs = idx[order, 0], vals[order]
    first = np.ones(n, bool)
    first[1:] = bs[1:] != bs[:-1]
    considered = first.copy()
    if preprocess_collapse_repeated:
        considered[1:] |= vs[1:] != vs[:-1]
    else:
        considered[:] = True
    null = considered & ((vs >= num_classes - 1) | (vs < 0))
    # number of terminators seen in this row BEFORE each entry
    c = np.cumsum(null)
    row_start = np.maximum.accumulate(np.where(first, np.arange(n), 0))
    before = c - null - (c[row_start] - null[row_start])
    emit = considered & ~null
    bad = emit & (before > 0)
    if bad.any():
        raise ValueError("ctc_loss: saw a non-null label (index >= num_classes - 1) following a null "
                         "label, batch: %d" % int(bs[np.argmax(bad)]))
    eb, ev = bs[emit], vs[emit]
    lens = np.bincount(eb, minlength=Bn).astype(np.int32)
    Lmax = max(int(lens.max(initial=0)), 1)
    lab = np.zeros((Bn, Lmax), dtype=np.int32)
    starts = np.concatenate([[0], np.cumsum(lens)[:-1]])
    pos = np.arange(eb.shape[0]) - starts[eb]
    lab[eb, pos] = ev
    reps = np.zeros(Bn, dtype=np.int32)
    if eb.shape[0] > 1:
        same = (eb[1:] == eb[:-1]) & (ev[1:] == ev[:-1])
        reps = np.bincount(eb[1:][same], minlength=Bn).astype(np.int32)
    return lab, lens, reps


def ctc_prepare_labels_device(indices: torch.Tensor, values: torch.Tensor, dense_shape, num_classes: int,
                              preprocess_collapse_repeated: bool):
    """Device form of ctc_prepare_labels: indices int64 [n,2] (row-major sorted) and values int32 [n]
    are CUDA tensors, nothing is copied to the host.  -> (labels [B,Lcap] int32, label_lens [B], repeats [B], status [B])
    CUDA tensors, Lcap = dense_shape[1]; status != 0 marks the rows for which tf.nn.ctc_loss would raise."""
    _need_cuda(indices, values)
    Bn, Lcap = int(dense_shape[0]), max(1, int(dense_shape[1]))
    idx = indices.contiguous().to(torch.int64)
    val = values.contiguous().to(torch.int32)
    dev = idx.device
    labels = torch.empty((Bn, Lcap), dtype=torch.int32, device=dev)
    lens = torch.empty((Bn,), dtype=torch.int32, device=dev)
    reps = torch.empty((Bn,), dtype=torch.int32, device=dev)
    status = torch.empty((Bn,), dtype=torch.int32, device=dev)
    check(lib.ds2_ctc_labels_from_sparse(_p(idx), _p(val), int(val.numel()), Bn, Lcap, num_classes, int(preprocess_collapse_repeated),
                                         _p(labels), _p(lens), _p(reps), _p(status), _stream()), "ds2_ctc_labels_from_sparse")
    return labels, lens, reps, status


class _CTCFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels_dev, label_lens_dev, seq_lens_dev, Lmax):
        T, B, K = logits.shape
        loss = torch.empty(B, dtype=torch.float32, device=logits.device)
        grad = torch.empty_like(logits)
        wsb = lib.ds2_ctc_workspace_bytes(T, B, K, Lmax)
        ws = _workspace(wsb, logits.device, "ctc")
        check(lib.ds2_ctc_loss_fwd_bwd(_p(logits), T, B, K, _p(labels_dev), Lmax, _p(label_lens_dev), _p(seq_lens_dev), None,
                                       _p(loss), _p(grad), _p(ws), wsb, _stream()), "ds2_ctc_loss_fwd_bwd")
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (grad,) = ctx.saved_tensors
        T, B, K = grad.shape
        out = torch.empty_like(grad)
        check(lib.ds2_ctc_scale_grad(_p(grad), _p(dloss.contiguous().float()), _p(out), T, B, K, _stream()), "ds2_ctc_scale_grad")
        return out, None, None, None, None


def ctc_loss_dense(logits: torch.Tensor, labels_np, label_lens_np, repeats_np, seq_lens) -> torch.Tensor:
    """Per-utterance CTC NLL [B] from packed labels; autograd-aware.  Raises ValueError where TF raises
    InvalidArgumentError ("Not enough time for target transition sequence")."""
    _need_cuda(logits)
    if logits.dtype != torch.float32:
        logits = cast(logits, torch.float32)             # CTC always runs in fp32 (src/deepSpeech_NCHW.py:225)
    T, B, K = logits.shape
    sl = np.asarray(seq_lens.detach().cpu().numpy() if isinstance(seq_lens, torch.Tensor) else seq_lens, dtype=np.int32)
    if sl.shape[0] != B or labels_np.shape[0] != B:
        raise ValueError("ctc_loss: batch size mismatch (logits %d, labels %d, sequence_length %d)" % (B, labels_np.shape[0], sl.shape[0]))
    if (sl > T).any() or (sl < 0).any():
        raise ValueError("ctc_loss: sequence_length must lie in [0, %d]" % T)
    need = label_lens_np + repeats_np
    bad = np.nonzero(sl < need)[0]
    if bad.size:
        b = int(bad[0])
        raise ValueError("ctc_loss: Not enough time for target transition sequence (required: %d, available: %d) batch: %d"
                         % (int(need[b]), int(sl[b]), b))
    if labels_np.size and (labels_np.max() >= K - 1 or labels_np.min() < 0):
        raise ValueError("ctc_loss: label out of range [0, %d)" % (K - 1))
    dev = logits.device
    host = np.concatenate([labels_np.reshape(-1), label_lens_np, sl]).astype(np.int32)
    packed = torch.from_numpy(host).to(dev, non_blocking=True)
    Lmax = labels_np.shape[1]
    lab_d = packed[:B * Lmax]
    ll_d = packed[B * Lmax:B * Lmax + B]
    sl_d = packed[B * Lmax + B:]
    return _CTCFn.apply(logits.contiguous(), lab_d, ll_d, sl_d, Lmax)


def ctc_loss_sparse_device(logits: torch.Tensor, indices: torch.Tensor, values: torch.Tensor, dense_shape, seq_lens,
                           preprocess_collapse_repeated: bool) -> torch.Tensor:
    """tf.nn.ctc_loss on SparseTensor labels that live on the device: label pre-processing (ds2_ctc_labels_from_sparse) and
    the loss without any device -> host transfer.  Rows that TF would reject carry label_len = -1 into the kernels, which
    answer loss = +inf and a zero gradient row (include/ds2_b200.h)."""
    _need_cuda(logits, indices, values)
    if logits.dtype != torch.float32:
        logits = cast(logits, torch.float32)
    T, B, K = logits.shape
    if int(dense_shape[0]) != B:
        raise ValueError("ctc_loss: batch size mismatch (logits %d, labels %d)" % (B, int(dense_shape[0])))
    lab, ll, reps, status = ctc_prepare_labels_device(indices, values, dense_shape, K, preprocess_collapse_repeated)
    if isinstance(seq_lens, torch.Tensor):
        sl = seq_lens.to(device=logits.device, dtype=torch.int32, non_blocking=True)
    else:
        sl = torch.from_numpy(np.ascontiguousarray(seq_lens, dtype=np.int32)).to(logits.device, non_blocking=True)
    if sl.shape[0] != B:
        raise ValueError("ctc_loss: batch size mismatch (logits %d, sequence_length %d)" % (B, sl.shape[0]))
    ll = torch.where(status != 0, torch.full_like(ll, -1), ll)
    return _CTCFn.apply(logits.contiguous(), lab, ll, sl, lab.shape[1])
