"""Tensor-level wrappers of the C ABI: torch only provides device memory and the stream.

Every function enqueues hand-written CUDA kernels from libb200ir.so on torch's current stream.
Activations are NCHW (possibly channel-slices of a wider buffer: only the batch stride is free).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib as L

lib = L.lib
_DT = {torch.float32: L.F32, torch.bfloat16: L.BF16}

# ---- instrumentation (bench.py): launch counter and optional per-kernel CUDA-event timing ----
LAUNCHES = 0            # kernels of libb200ir.so enqueued since import (each wrapper adds its launch count)
_TIMING = None          # None, or dict: key -> list of (start_event, end_event)
_TIMING_FILTER = None   # None = time every op; else a set of op names


def timing_start(only=None):
    global _TIMING, _TIMING_FILTER
    _TIMING, _TIMING_FILTER = {}, (set(only) if only else None)


def timing_stop():
    """-> {key: (n_launches, mean_ms)}; caller must have synchronised the stream."""
    global _TIMING
    t, _TIMING = _TIMING, None
    return {k: (len(v), sum(a.elapsed_time(b) for a, b in v) / len(v)) for k, v in (t or {}).items()}


class _timed:
    __slots__ = ("key", "ev")

    def __init__(self, name, *shape):
        global LAUNCHES
        LAUNCHES += 1
        self.key = None
        if _TIMING is not None and (_TIMING_FILTER is None or name in _TIMING_FILTER):
            self.key = (name,) + tuple(shape)

    def __enter__(self):
        if self.key is not None:
            self.ev = torch.cuda.Event(enable_timing=True)
            self.ev.record()
        return self

    def __exit__(self, *a):
        if self.key is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            _TIMING.setdefault(self.key, []).append((self.ev, e))
        return False


# torch.cuda.current_stream() / current_device() resolve the device through several Python layers (~10 us per call under
# load: a quarter of the host time of an eager training step); the raw getters are one C call each
_cur_device = getattr(torch._C, "_cuda_getDevice", None) or torch.cuda.current_device
_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream() -> int:
    if _raw_stream is not None:
        return _raw_stream(_cur_device())
    return torch.cuda.current_stream().cuda_stream


def _act(t: torch.Tensor, name: str):
    """(ptr, batch_stride, dtype_code) of an NCHW activation view."""
    if not t.is_cuda:
        raise ValueError(f"{name}: expected a CUDA tensor (this package has no CPU fallback)")
    if t.device.index != _cur_device():
        # every entry point enqueues on the CURRENT device's stream and sets per-device kernel attributes for it: a tensor
        # of another GPU would be dereferenced on the wrong device
        raise ValueError(f"{name}: tensor lives on cuda:{t.device.index} but the current device is "
                         f"cuda:{torch.cuda.current_device()} (use torch.cuda.set_device / torch.cuda.device_of(x))")
    dt = _DT.get(t.dtype)
    if dt is None:
        raise ValueError(f"{name}: dtype {t.dtype} unsupported (float32 / bfloat16)")
    if t.dim() != 4:
        raise ValueError(f"{name}: expected NCHW, got shape {tuple(t.shape)}")
    if t.is_contiguous():      # the common case, one C call instead of the stride arithmetic below
        sh = t.shape
        return t.data_ptr(), sh[1] * sh[2] * sh[3], dt
    b, c, h, w = t.shape
    s = t.stride()
    ok = (s[3] == 1 or w == 1) and (s[2] == w or h == 1) and (s[1] == h * w or c == 1)
    if not ok:
        raise ValueError(f"{name}: not NCHW-contiguous within a sample (shape {tuple(t.shape)}, strides {s})")
    bs = s[0] if b > 1 else c * h * w
    return t.data_ptr(), bs, _DT[t.dtype]


def _par(t: Optional[torch.Tensor], name: str, numel: Optional[int] = None) -> Optional[int]:
    """Device pointer of an fp32 contiguous parameter (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
        raise ValueError(f"{name}: parameters must be contiguous float32 CUDA tensors")
    if numel is not None and t.numel() != numel:
        raise ValueError(f"{name}: expected {numel} elements, got {t.numel()}")
    return t.data_ptr()


def ln_conv1x1(x, w, bias=None, ln_type=L.LN_NONE, ln_w=None, ln_b=None, res=None, out=None,
               out_dtype=None, per_sample_w=False):
    """y = res + bias + W . LN(x)   (LayerNorm + 1x1 conv + residual), fp32 CUDA-core path."""
    B, Cin, H, W_ = x.shape
    Cout = w.shape[-2] if per_sample_w else w.shape[0]
    xp, xbs, xdt = _act(x, "x")
    if out is None:
        out = torch.empty((B, Cout, H, W_), device=x.device, dtype=out_dtype or x.dtype)
    yp, ybs, ydt = _act(out, "out")
    rp, rbs, rdt = (None, 0, 0)
    if res is not None:
        if tuple(res.shape) != (B, Cout, H, W_):
            raise ValueError("res shape mismatch")
        rp, rbs, rdt = _act(res, "res")
    wp = _par(w, "w", (B if per_sample_w else 1) * Cout * Cin)
    with _timed("ln_conv1x1", B, Cin, Cout, H * W_, ln_type, int(res is not None), int(per_sample_w), xdt, ydt):
      rc = lib.b200ir_ln_conv1x1_fwd(xp, xbs, xdt, wp, Cout * Cin if per_sample_w else 0, _par(bias, "bias", Cout),
                                   ln_type, _par(ln_w, "ln_w", Cin if ln_type else None),
                                   _par(ln_b, "ln_b", Cin if ln_type == L.LN_WITHBIAS else None),
                                     rp, rbs, rdt, yp, ybs, ydt, B, Cin, Cout, H * W_, _stream())
    L.check(rc, "b200ir_ln_conv1x1_fwd")
    return out


def dwconv3x3(x, w, bias=None, out=None):
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if out is None:
        out = torch.empty_like(x, memory_format=torch.contiguous_format)
    yp, ybs, ydt = _act(out, "out")
    if ydt != xdt:
        raise ValueError("dwconv3x3: in/out dtype mismatch")
    with _timed("dwconv3x3", B, C, H, W_, xdt):
        rc = lib.b200ir_dwconv3x3_fwd(xp, xbs, _par(w, "w", C * 9), _par(bias, "bias", C), yp, ybs, xdt, B, C, H, W_,
                                      _stream())
    L.check(rc, "b200ir_dwconv3x3_fwd")
    return out


def dwconv_kxk(x, w, bias=None, out=None):
    """Depthwise K x K conv (K = 3, 5, 7; stride 1, zero padding K//2)."""
    B, C, H, W_ = x.shape
    K = w.shape[-1]
    xp, xbs, xdt = _act(x, "x")
    if out is None:
        out = torch.empty_like(x, memory_format=torch.contiguous_format)
    yp, ybs, ydt = _act(out, "out")
    if ydt != xdt:
        raise ValueError("dwconv_kxk: in/out dtype mismatch")
    with _timed("dwconv_kxk", B, C, H, W_, K, xdt):
        rc = lib.b200ir_dwconv_kxk_fwd(xp, xbs, _par(w, "w", C * K * K), _par(bias, "bias", C), yp, ybs, xdt, B, C, H, W_,
                                       K, _stream())
    L.check(rc, "b200ir_dwconv_kxk_fwd")
    return out


def dwconv_kxk_bwd(dy, x, w, want_bias=False, want_input=True):
    """Backward of dwconv_kxk: (dx or None, dw [C,1,K,K], db [C] or None)."""
    B, C, H, W_ = dy.shape
    K = w.shape[-1]
    gp, gbs, gdt = _act(dy, "dy")
    xp, xbs, xdt = _act(x, "x")
    if gdt != xdt:
        raise ValueError("dwconv_kxk_bwd: dtype mismatch")
    dx = None
    if want_input:
        dx = torch.empty_like(dy, memory_format=torch.contiguous_format)
        dp, dbs, _ = _act(dx, "dx")
        with _timed("dwconv_kxk", B, C, H, W_, K, gdt):
            rc = lib.b200ir_dwconv_kxk_bwd_input(gp, gbs, _par(w, "w", C * K * K), dp, dbs, gdt, B, C, H, W_, K, _stream())
        L.check(rc, "b200ir_dwconv_kxk_bwd_input")
    nparts = lib.b200ir_dwconv_kxk_wgrad_nparts(B, H, W_)
    nv = K * K + 1
    parts = torch.empty((nparts, C, nv), device=dy.device, dtype=torch.float32)
    with _timed("dwconv_kxk_wgrad", B, C, H, W_, K, gdt):
        rc = lib.b200ir_dwconv_kxk_wgrad(gp, gbs, xp, xbs, gdt, parts.data_ptr(), nparts, B, C, H, W_, K, _stream())
    L.check(rc, "b200ir_dwconv_kxk_wgrad")
    dw, db = reduce_partials_taps(parts, nparts, C, nv, want_bias)
    return dx, dw.view(C, 1, K, K), db


def sigmoid_gate(u, g, out=None):
    """y = u * sigmoid(g); g has G channels, each gating C/G consecutive channels of u (G = C or G = heads)."""
    B, C, H, W_ = u.shape
    G = g.shape[1]
    up, ubs, udt = _act(u, "u")
    gp, gbs, gdt = _act(g, "g")
    if gdt != udt or g.shape[0] != B or tuple(g.shape[2:]) != (H, W_):
        raise ValueError("sigmoid_gate: u / g mismatch")
    if out is None:
        out = torch.empty((B, C, H, W_), device=u.device, dtype=u.dtype)
    yp, ybs, _ = _act(out, "out")
    with _timed("sigmoid_gate", B, C, G, H * W_, udt):
        rc = lib.b200ir_sigmoid_gate_fwd(up, ubs, gp, gbs, udt, yp, ybs, B, C, G, H * W_, _stream())
    L.check(rc, "b200ir_sigmoid_gate_fwd")
    return out


def sigmoid_gate_bwd(u, g, dy):
    """-> (du, dg)."""
    B, C, H, W_ = u.shape
    G = g.shape[1]
    up, ubs, udt = _act(u, "u")
    gp, gbs, _ = _act(g, "g")
    dp, dbs, ddt = _act(dy, "dy")
    if ddt != udt:
        raise ValueError("sigmoid_gate_bwd: dtype mismatch")
    du = torch.empty((B, C, H, W_), device=u.device, dtype=u.dtype)
    dg = torch.empty((B, G, H, W_), device=u.device, dtype=u.dtype)
    dup, dubs, _ = _act(du, "du")
    dgp, dgbs, _ = _act(dg, "dg")
    with _timed("sigmoid_gate_bwd", B, C, G, H * W_, udt):
        rc = lib.b200ir_sigmoid_gate_bwd(up, ubs, gp, gbs, dp, dbs, udt, dup, dubs, dgp, dgbs, B, C, G, H * W_, _stream())
    L.check(rc, "b200ir_sigmoid_gate_bwd")
    return du, dg


def dwconv3x3_bwd_input(dy, w):
    """Input gradient of the depthwise 3x3 conv: depthwise conv of dy with w's taps in reverse order."""
    B, C, H, W_ = dy.shape
    gp, gbs, gdt = _act(dy, "dy")
    dx = torch.empty_like(dy, memory_format=torch.contiguous_format)
    dp, dbs, _ = _act(dx, "dx")
    with _timed("dwconv3x3", B, C, H, W_, gdt):
        rc = lib.b200ir_dwconv3x3_bwd_input(gp, gbs, _par(w, "w", C * 9), dp, dbs, gdt, B, C, H, W_, _stream())
    L.check(rc, "b200ir_dwconv3x3_bwd_input")
    return dx


def dwconv3x3_gelu_gate(x, w, bias=None, out=None):
    B, C2, H, W_ = x.shape
    if C2 % 2:
        raise ValueError("dwconv3x3_gelu_gate: channel count must be even")
    Hd = C2 // 2
    xp, xbs, xdt = _act(x, "x")
    if out is None:
        out = torch.empty((B, Hd, H, W_), device=x.device, dtype=x.dtype)
    yp, ybs, ydt = _act(out, "out")
    if ydt != xdt:
        raise ValueError("dwconv3x3_gelu_gate: in/out dtype mismatch")
    with _timed("dwconv3x3_gelu_gate", B, Hd, H, W_, xdt):
        rc = lib.b200ir_dwconv3x3_gelu_gate_fwd(xp, xbs, _par(w, "w", C2 * 9), _par(bias, "bias", C2), yp, ybs, xdt,
                                                B, Hd, H, W_, _stream())
    L.check(rc, "b200ir_dwconv3x3_gelu_gate_fwd")
    return out


def dwconv3x3_gelu_sigmoid_gate(x, w, bias=None, out=None):
    """out[c] = gelu(dw(x)[c]) * sigmoid(dw(x)[c + Hd]) (split-gate FFN, experiment/RectiFormer.py:231-266)."""
    B, C2, H, W_ = x.shape
    Hd = C2 // 2
    xp, xbs, xdt = _act(x, "x")
    if out is None:
        out = torch.empty((B, Hd, H, W_), device=x.device, dtype=x.dtype)
    yp, ybs, ydt = _act(out, "out")
    if ydt != xdt or C2 % 2:
        raise ValueError("dwconv3x3_gelu_sigmoid_gate: dtype / channel mismatch")
    with _timed("dwconv3x3_gelu_sigmoid_gate", B, Hd, H, W_, xdt):
        rc = lib.b200ir_dwconv3x3_gelu_sigmoid_gate_fwd(xp, xbs, _par(w, "w", C2 * 9), _par(bias, "bias", C2), yp, ybs,
                                                        xdt, B, Hd, H, W_, _stream())
    L.check(rc, "b200ir_dwconv3x3_gelu_sigmoid_gate_fwd")
    return out


def mdta_gram(qkv, heads, impl=L.IMPL_AUTO):
    """Partial Grams / norms per (b, head, HW-slab): fp32 [B, heads, nslab, d*d+2d]."""
    B, C3, H, W_ = qkv.shape
    C = C3 // 3
    d = C // heads
    qp, qbs, qdt = _act(qkv, "qkv")
    nslab = lib.b200ir_mdta_gram_nslab(B, C, heads, H * W_)
    if nslab <= 0:
        raise RuntimeError("b200ir_mdta_gram_nslab failed")
    partials = torch.empty((B, heads, nslab, d * d + 2 * d), device=qkv.device, dtype=torch.float32)
    with _timed("mdta_gram", B, C, heads, H * W_, qdt):
        rc = lib.b200ir_mdta_gram_fwd(qp, qbs, qdt, B, C, heads, H * W_, partials.data_ptr(), nslab, impl, _stream())
    L.check(rc, "b200ir_mdta_gram_fwd")
    return partials


def mdta_softmax_fold(partials, temperature, w_out, want_attn=False, packed=False):
    """attn = softmax(normalised Gram * temperature); m = W_out . blockdiag(attn).
    packed=False -> fp32 m [B,C,C]; packed=True -> bf16 [B,nrows,kpad] operand of gemm_tc.  (+attn if asked)"""
    B, heads, nslab, per = partials.shape
    C = w_out.shape[0]
    d = C // heads
    attn = torch.empty((B, heads, d, d), device=partials.device, dtype=torch.float32)
    m, mp, kpad, nrows = None, None, 0, 0
    if packed:
        kpad, nrows = gemm_tc_geometry(C, C)
        mp = torch.empty((B, nrows, kpad), device=partials.device, dtype=torch.bfloat16)
    else:
        m = torch.empty((B, C, C), device=partials.device, dtype=torch.float32)
    global LAUNCHES
    LAUNCHES += 1  # two kernels (softmax, fold)
    per_sample = temperature.numel() == B * heads and temperature.dim() == 2     # [B, heads]: dynamic temperature
    with _timed("mdta_softmax_fold", B, C, heads, nslab):
        rc = lib.b200ir_mdta_softmax_fold_fwd_ex(partials.data_ptr(), nslab,
                                                 _par(temperature, "temperature", B * heads if per_sample else heads),
                                                 int(per_sample), _par(w_out, "w_out", C * C),
                                                 m.data_ptr() if m is not None else None, attn.data_ptr(),
                                                 mp.data_ptr() if mp is not None else None, kpad, nrows, B, C, heads,
                                                 _stream())
    L.check(rc, "b200ir_mdta_softmax_fold_fwd")
    out = mp if packed else m
    return (out, attn) if want_attn else out


def conv3x3(x, w, bias=None, res=None, shuffle=L.SHUFFLE_NONE, out=None, out_dtype=None):
    B, Cin, H, W_ = x.shape
    Cout = w.shape[0]
    xp, xbs, xdt = _act(x, "x")
    if shuffle == L.UNSHUFFLE2:
        oshape = (B, Cout * 4, H // 2, W_ // 2)
    elif shuffle == L.SHUFFLE2:
        oshape = (B, Cout // 4, H * 2, W_ * 2)
    else:
        oshape = (B, Cout, H, W_)
    if out is None:
        out = torch.empty(oshape, device=x.device, dtype=out_dtype or x.dtype)
    elif tuple(out.shape) != oshape:
        raise ValueError(f"conv3x3: out shape {tuple(out.shape)} != {oshape}")
    yp, ybs, ydt = _act(out, "out")
    rp, rbs, rdt = (None, 0, 0)
    if res is not None:
        if tuple(res.shape) != oshape:
            raise ValueError("conv3x3: res shape mismatch")
        rp, rbs, rdt = _act(res, "res")
    with _timed("conv3x3", B, Cin, Cout, H, W_, shuffle, xdt, ydt):
        rc = lib.b200ir_conv3x3_fwd(xp, xbs, xdt, _par(w, "w", Cout * Cin * 9), _par(bias, "bias", Cout), rp, rbs, rdt,
                                    yp, ybs, ydt, B, Cin, Cout, H, W_, shuffle, _stream())
    L.check(rc, "b200ir_conv3x3_fwd")
    return out


def copy_cast(x, out):
    """out[...] = x (dtype cast allowed); both NCHW views of identical shape."""
    if tuple(x.shape) != tuple(out.shape):
        raise ValueError("copy_cast: shape mismatch")
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    yp, ybs, ydt = _act(out, "out")
    with _timed("copy_cast", B, C * H * W_, xdt, ydt):
        L.check(lib.b200ir_copy_cast(xp, xbs, xdt, yp, ybs, ydt, B, C * H * W_, _stream()), "b200ir_copy_cast")
    return out


# ---------------------------------------------------------------------------------------------------
# tcgen05 path
# ---------------------------------------------------------------------------------------------------
import ctypes as _C

_GEOM = {}


def gemm_tc_geometry(Cin, Cout):
    """(kpad, nrows) of the packed weight, or None if the tcgen05 GEMM does not support (Cin, Cout)."""
    key = (Cin, Cout)
    if key not in _GEOM:
        kp, nr = _C.c_int(0), _C.c_int(0)
        rc = lib.b200ir_gemm_tc_geometry(Cin, Cout, _C.byref(kp), _C.byref(nr))
        _GEOM[key] = (kp.value, nr.value) if rc == 0 else None
    return _GEOM[key]


def gemm_tc_usable(x, Cout):
    B, Cin, H, W_ = x.shape
    hw = H * W_
    return (x.dtype == torch.bfloat16 and hw % 8 == 0 and hw >= 64 and x.data_ptr() % 16 == 0
            and (B == 1 or x.stride(0) % 8 == 0) and gemm_tc_geometry(Cin, Cout) is not None)


def pack_conv1x1(w, bias=None, ln_type=L.LN_NONE, ln_w=None, ln_b=None, per_sample=False):
    """Pack fp32 1x1-conv weights ([Cout,Cin] or [B,Cout,Cin]) for gemm_tc: (wp bf16 [Bw,nrows,kpad], st fp32 [2,nrows])."""
    if per_sample:
        Bw, Cout, Cin = w.shape
    else:
        Bw, Cout, Cin = 1, w.shape[0], w.shape[1]
    geo = gemm_tc_geometry(Cin, Cout)
    if geo is None:
        raise ValueError(f"pack_conv1x1: ({Cin},{Cout}) unsupported by the tcgen05 GEMM")
    kpad, nrows = geo
    wp = torch.empty((Bw, nrows, kpad), device=w.device, dtype=torch.bfloat16)
    st = torch.empty((2, nrows), device=w.device, dtype=torch.float32)
    with _timed("pack_conv1x1", Bw, Cin, Cout):
        rc = lib.b200ir_pack_conv1x1(_par(w, "w", Bw * Cout * Cin), Cout * Cin if per_sample else 0, Bw,
                                     _par(bias, "bias", Cout), ln_type, _par(ln_w, "ln_w", Cin if ln_type else None),
                                     _par(ln_b, "ln_b", Cin if ln_type == L.LN_WITHBIAS else None), Cin, Cout,
                                     wp.data_ptr(), st.data_ptr(), _stream())
    L.check(rc, "b200ir_pack_conv1x1")
    return wp, st


def pack_conv1x1_multi(descs_dev, ndesc, total_blocks):
    """One launch for a whole table of packs (descs_dev: device uint8 tensor holding b200ir_pack_desc_t[ndesc])."""
    with _timed("pack_conv1x1_multi", ndesc, total_blocks):
        rc = lib.b200ir_pack_conv1x1_multi(descs_dev.data_ptr(), ndesc, total_blocks, _stream())
    L.check(rc, "b200ir_pack_conv1x1_multi")


def pack_conv1x1_t(w):
    """Operand of the input-gradient GEMM dx = W^T.dy: same result as pack_conv1x1(w.t()), packed from w [Cout,Cin]."""
    Cout, Cin = w.shape
    geo = gemm_tc_geometry(Cout, Cin)
    if geo is None:
        raise ValueError(f"pack_conv1x1_t: ({Cout},{Cin}) unsupported by the tcgen05 GEMM")
    kpad, nrows = geo
    wp = torch.empty((1, nrows, kpad), device=w.device, dtype=torch.bfloat16)
    st = torch.empty((2, nrows), device=w.device, dtype=torch.float32)
    with _timed("pack_conv1x1", 1, Cout, Cin):
        rc = lib.b200ir_pack_conv1x1_t(_par(w, "w", Cout * Cin), Cin, Cout, wp.data_ptr(), st.data_ptr(), _stream())
    L.check(rc, "b200ir_pack_conv1x1_t")
    return wp, st


def gemm_tc(x, packed, Cout, ln_type=L.LN_NONE, res=None, out=None, out_dtype=None, variant=L.GEMM_AUTO):
    """y = res + W.LN(x) + bias on tcgen05 tensor cores; `packed` = pack_conv1x1(...).  `variant` forces a kernel
    (L.GEMM_RESIDENT / GEMM_STREAM / GEMM_RESIDENT_CH) instead of the shape-based choice (tests, A/B timing)."""
    wp, st = packed
    B, Cin, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if xdt != L.BF16:
        raise ValueError("gemm_tc: activations must be bfloat16")
    if out is None:
        out = torch.empty((B, Cout, H, W_), device=x.device, dtype=out_dtype or x.dtype)
    yp, ybs, ydt = _act(out, "out")
    rp, rbs, rdt = (None, 0, 0)
    if res is not None:
        if tuple(res.shape) != (B, Cout, H, W_):
            raise ValueError("res shape mismatch")
        rp, rbs, rdt = _act(res, "res")
    Bw = wp.shape[0]
    with _timed("gemm_tc", B, Cin, Cout, H * W_, ln_type, int(res is not None), int(Bw > 1), xdt, ydt):
        rc = lib.b200ir_gemm_tc_fwd_ex(xp, xbs, wp.data_ptr(), Bw, st.data_ptr(), ln_type, rp, rbs, rdt, yp, ybs, ydt,
                                       B, Cin, Cout, H * W_, variant, _stream())
    L.check(rc, "b200ir_gemm_tc_fwd")
    return out


_GEOMF = {}


def fused_proj_dw_geometry(C, Cout, gate):
    """(kf, nrows) of the packed weight of the fused LN + 1x1 + depthwise kernel, or None if (C, Cout) is unsupported."""
    key = (C, Cout, bool(gate))
    if key not in _GEOMF:
        kf, nr = _C.c_int(0), _C.c_int(0)
        rc = lib.b200ir_fused_proj_dw_geometry(C, Cout, int(bool(gate)), _C.byref(kf), _C.byref(nr))
        _GEOMF[key] = (kf.value, nr.value) if rc == 0 else None
    return _GEOMF[key]


def ln_proj_dw_usable(x, Cout, gate):
    B, C, H, W_ = x.shape
    return (x.dtype == torch.bfloat16 and x.data_ptr() % 16 == 0 and (B == 1 or x.stride(0) % 8 == 0)
            and lib.b200ir_ln_proj_dw_supported(C, Cout, H, W_, int(bool(gate))) == 1)


def pack_fused_proj_dw(w, bias, ln_type, ln_w, ln_b, Cout, gate):
    """Pack the fp32 1x1 weight [Ctot, C] (Ctot = 2*Cout with the gate, Cout without) for ln_proj_dw: bf16 [nrows, kf]."""
    Ctot, Cin = w.shape[0], w.shape[1]
    if Ctot != (2 * Cout if gate else Cout):
        raise ValueError("pack_fused_proj_dw: weight rows do not match Cout / gate")
    geo = fused_proj_dw_geometry(Cin, Cout, gate)
    if geo is None:
        raise ValueError(f"pack_fused_proj_dw: ({Cin},{Cout}) unsupported")
    kf, nrows = geo
    wf = torch.empty((nrows, kf), device=w.device, dtype=torch.bfloat16)
    with _timed("pack_fused_proj_dw", Cin, Cout, int(bool(gate))):
        rc = lib.b200ir_pack_fused_proj_dw(_par(w, "w", Ctot * Cin), _par(bias, "bias", Ctot), ln_type,
                                           _par(ln_w, "ln_w", Cin if ln_type else None),
                                           _par(ln_b, "ln_b", Cin if ln_type == L.LN_WITHBIAS else None), Cin, Cout,
                                           int(bool(gate)), wf.data_ptr(), _stream())
    L.check(rc, "b200ir_pack_fused_proj_dw")
    return wf


def ln_proj_dw(x, wf, dw_w, dw_b, ln_type, Cout, gate, out=None):
    """out = [gelu-gate of] dwconv3x3(conv1x1(LN(x))) in ONE kernel; `wf` = pack_fused_proj_dw(...)."""
    B, Cin, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if xdt != L.BF16:
        raise ValueError("ln_proj_dw: activations must be bfloat16")
    if out is None:
        out = torch.empty((B, Cout, H, W_), device=x.device, dtype=torch.bfloat16)
    yp, ybs, ydt = _act(out, "out")
    if ydt != L.BF16 or tuple(out.shape) != (B, Cout, H, W_):
        raise ValueError("ln_proj_dw: out must be bfloat16 [B, Cout, H, W]")
    Ctot = 2 * Cout if gate else Cout
    with _timed("ln_proj_dw", B, Cin, Cout, H, W_, int(bool(gate))):
        rc = lib.b200ir_ln_proj_dw_fwd(xp, xbs, wf.data_ptr(), _par(dw_w, "dw_w", Ctot * 9), _par(dw_b, "dw_b", Ctot),
                                       ln_type, yp, ybs, B, Cin, Cout, H, W_, int(bool(gate)), _stream())
    L.check(rc, "b200ir_ln_proj_dw_fwd")
    return out


_GEOM3 = {}


def conv3x3_tc_geometry(Cin, Cout):
    key = (Cin, Cout)
    if key not in _GEOM3:
        kp, nr = _C.c_int(0), _C.c_int(0)
        rc = lib.b200ir_conv3x3_tc_geometry(Cin, Cout, _C.byref(kp), _C.byref(nr))
        _GEOM3[key] = (kp.value, nr.value) if rc == 0 else None
    return _GEOM3[key]


def conv3x3_tc_usable(x, Cout):
    B, Cin, H, W_ = x.shape
    return (x.dtype == torch.bfloat16 and Cin % 8 == 0 and Cout >= 16
            and conv3x3_tc_geometry(Cin, Cout) is not None)


def nchw_to_nhwc(x):
    """bf16 [B,C,H,W] -> [B,H,W,C] staging copy (activation operand of the tcgen05 3x3 conv)."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if xdt != L.BF16:
        raise ValueError("nchw_to_nhwc: bfloat16 only")
    y = torch.empty((B, H, W_, C), device=x.device, dtype=torch.bfloat16)
    with _timed("nchw_to_nhwc", B, C, H * W_):
        rc = lib.b200ir_nchw_to_nhwc(xp, xbs, y.data_ptr(), B, C, H * W_, _stream())
    L.check(rc, "b200ir_nchw_to_nhwc")
    return y


def pack_conv3x3(w, bias=None):
    Cout, Cin = w.shape[0], w.shape[1]
    geo = conv3x3_tc_geometry(Cin, Cout)
    if geo is None:
        raise ValueError(f"pack_conv3x3: ({Cin},{Cout}) unsupported by the tcgen05 conv")
    kpad, nrows = geo
    wp = torch.empty((nrows, kpad), device=w.device, dtype=torch.bfloat16)
    st = torch.empty((2, nrows), device=w.device, dtype=torch.float32)
    with _timed("pack_conv3x3", Cin, Cout):
        rc = lib.b200ir_pack_conv3x3(_par(w, "w", Cout * Cin * 9), _par(bias, "bias", Cout), Cin, Cout, wp.data_ptr(),
                                     st.data_ptr(), _stream())
    L.check(rc, "b200ir_pack_conv3x3")
    return wp, st


def conv3x3_tc(x, packed, Cout, shuffle=L.SHUFFLE_NONE, out=None, out_dtype=None):
    wp, st = packed
    B, Cin, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if xdt != L.BF16:
        raise ValueError("conv3x3_tc: activations must be bfloat16")
    if shuffle == L.UNSHUFFLE2:
        oshape = (B, Cout * 4, H // 2, W_ // 2)
    elif shuffle == L.SHUFFLE2:
        oshape = (B, Cout // 4, H * 2, W_ * 2)
    else:
        oshape = (B, Cout, H, W_)
    if out is None:
        out = torch.empty(oshape, device=x.device, dtype=out_dtype or x.dtype)
    elif tuple(out.shape) != oshape:
        raise ValueError(f"conv3x3_tc: out shape {tuple(out.shape)} != {oshape}")
    yp, ybs, ydt = _act(out, "out")
    xt = nchw_to_nhwc(x)
    with _timed("conv3x3_tc", B, Cin, Cout, H, W_, shuffle, xdt, ydt):
        rc = lib.b200ir_conv3x3_tc_fwd(xt.data_ptr(), H * W_ * Cin, wp.data_ptr(), st.data_ptr(), yp, ybs, ydt, B, Cin,
                                       Cout, H, W_, shuffle, _stream())
    L.check(rc, "b200ir_conv3x3_tc_fwd")
    return out


# ---------------------------------------------------------------------------------------------------
# backward / training-step kernels (csrc/backward.cu)
# ---------------------------------------------------------------------------------------------------
def reduce_partials(parts, n, K, stride=None, out=None):
    """out[k] = sum_{i<n} parts.flat[i*stride + k]  (fixed order; fp32)."""
    if out is None:
        out = torch.empty((K,), device=parts.device, dtype=torch.float32)
    with _timed("reduce_partials", n, K):
        L.check(lib.b200ir_reduce_partials(parts.data_ptr(), out.data_ptr(), n, K, K if stride is None else stride,
                                           _stream()), "b200ir_reduce_partials")
    return out


def reduce_partials_taps(parts, nparts, C, nv, want_bias):
    """(dw [C, nv-1], db [C] or None) = fold of the depthwise partial sums parts [nparts, C, nv]: dense tensors (a strided
    view as a parameter gradient would push the optimizer's gradient gather off torch's multi-tensor route)."""
    dw = torch.empty((C, nv - 1), device=parts.device, dtype=torch.float32)
    db = torch.empty((C,), device=parts.device, dtype=torch.float32) if want_bias else None
    with _timed("reduce_partials", nparts, C * nv):
        L.check(lib.b200ir_reduce_partials_taps(parts.data_ptr(), dw.data_ptr(), db.data_ptr() if want_bias else None, nparts, C,
                                                nv, _stream()), "b200ir_reduce_partials_taps")
    return dw, db


def ln_fwd(x, ln_type, ln_w, ln_b, out_dtype=None):
    """LN(x) materialised (only the backward needs it as a tensor)."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    y = torch.empty((B, C, H, W_), device=x.device, dtype=out_dtype or x.dtype)
    yp, ybs, ydt = _act(y, "y")
    with _timed("ln_fwd", B, C, H * W_, xdt):
        rc = lib.b200ir_ln_fwd(xp, xbs, xdt, ln_type, _par(ln_w, "ln_w", C),
                               _par(ln_b, "ln_b", C if ln_type == L.LN_WITHBIAS else None), yp, ybs, ydt, B, C, H * W_,
                               _stream())
    L.check(rc, "b200ir_ln_fwd")
    return y


def ln_bwd(x, g, ln_type, ln_w, res=None, ln_b=None, want_xn=False, fold=None):
    """-> (dx, dweight[C], dbias[C]) or, with want_xn, (dx, dweight, dbias, LN(x)): the normalised tensor is a by-product
    of the same pass (operand of the following weight-gradient GEMM)."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    gp, gbs, gdt = _act(g, "g")
    dx = torch.empty((B, C, H, W_), device=x.device, dtype=x.dtype)
    dp, dbs, ddt = _act(dx, "dx")
    rp, rbs, rdt = (None, 0, 0)
    if res is not None:
        rp, rbs, rdt = _act(res, "res")
    xn, np_, nbs = None, None, 0
    if want_xn:
        xn = torch.empty((B, C, H, W_), device=x.device, dtype=x.dtype)
        np_, nbs, _ = _act(xn, "xn")
    nparts = lib.b200ir_ln_bwd_nparts(B, H * W_)
    parts = torch.empty((nparts, 2, C), device=x.device, dtype=torch.float32)
    with _timed("ln_bwd", B, C, H * W_, xdt):
        rc = lib.b200ir_ln_bwd(xp, xbs, xdt, gp, gbs, gdt, ln_type, _par(ln_w, "ln_w", C),
                               _par(ln_b, "ln_b", C) if ln_type == L.LN_WITHBIAS and ln_b is not None else None,
                               rp, rbs, rdt, dp, dbs, ddt, np_, nbs, parts.data_ptr(), nparts, B, C, H * W_, _stream())
    L.check(rc, "b200ir_ln_bwd")
    # `fold`: a callable that runs the fold of the (dweight | dbias) partial sums wherever the caller wants it (the
    # training step puts it on its side stream); default = here, on the current stream
    d = fold(lambda: reduce_partials(parts, nparts, 2 * C), parts) if fold is not None else reduce_partials(parts, nparts, 2 * C)
    if want_xn:
        return dx, d[:C], d[C:], xn
    return dx, d[:C], d[C:]


def conv_wgrad(dy, x, taps=1, per_sample=False, want_bias=False, impl=L.IMPL_AUTO):
    """Weight (and bias) gradient of a 1x1 / dense 3x3 conv: (dW [G,Cout,Cin*taps] or [Cout,Cin*taps], db or None)."""
    B, Cout, H, W_ = dy.shape
    Cin = x.shape[1]
    if x.shape[0] != B or tuple(x.shape[2:]) != (H, W_):
        raise ValueError("conv_wgrad: dy / x shape mismatch")
    if (taps == 9 and impl != L.IMPL_SIMT and torch.bfloat16 in (dy.dtype, x.dtype) and (H * W_) % 8 == 0
            and H * W_ >= 64 and B * Cin * 9 * H * W_ * 2 <= (4 << 30)):
        # bf16 training path: im2col (bf16) + the tcgen05 GEMM over Cin*9 "channels"
        x9 = torch.empty((B, Cin * 9, H, W_), device=x.device, dtype=torch.bfloat16)
        xp, xbs, xdt = _act(x, "x")
        with _timed("im2col3x3", B, Cin, H, W_, xdt):
            L.check(lib.b200ir_im2col3x3(xp, xbs, xdt, x9.data_ptr(), B, Cin, H, W_, _stream()), "b200ir_im2col3x3")
        if dy.dtype != torch.bfloat16:
            dy = copy_cast(dy, torch.empty(dy.shape, device=dy.device, dtype=torch.bfloat16))
        return conv_wgrad(dy, x9, 1, per_sample, want_bias, impl)
    dp, dbs, ddt = _act(dy, "dy")
    xp, xbs, xdt = _act(x, "x")
    ns = lib.b200ir_conv_wgrad_nsplit(B, Cout, Cin, taps, H * W_, int(per_sample))
    if ns <= 0:
        raise RuntimeError("b200ir_conv_wgrad_nsplit failed")
    G = B if per_sample else 1
    N = Cin * taps
    pw = torch.empty((ns, G, Cout, N), device=dy.device, dtype=torch.float32)
    pb = torch.empty((ns, G, Cout), device=dy.device, dtype=torch.float32) if want_bias else None
    with _timed("conv_wgrad", B, Cout, Cin, H * W_, taps, int(per_sample), ddt, xdt):
        rc = lib.b200ir_conv_wgrad(dp, dbs, ddt, xp, xbs, xdt, pw.data_ptr(), pb.data_ptr() if want_bias else None, ns,
                                   B, Cout, Cin, H, W_, taps, int(per_sample), impl, _stream())
    L.check(rc, "b200ir_conv_wgrad")
    dw = pw[0] if ns == 1 else reduce_partials(pw, ns, G * Cout * N).view(G, Cout, N)
    db = None
    if want_bias:  # the bias of a per-sample weight is still shared: sum over samples and splits
        db = pb.view(Cout) if ns * G == 1 else reduce_partials(pb, ns * G, Cout)
    return (dw if per_sample else dw[0]), db


def dwconv3x3_wgrad(dy, x, want_bias=False):
    """-> (dw [C,1,3,3], db [C] or None)."""
    B, C, H, W_ = x.shape
    dp, dbs, ddt = _act(dy, "dy")
    xp, xbs, xdt = _act(x, "x")
    if ddt != xdt:
        raise ValueError("dwconv3x3_wgrad: dtype mismatch")
    nparts = lib.b200ir_dwconv3x3_bwd_nparts(B, C, H, W_)
    parts = torch.empty((nparts, C, 10), device=x.device, dtype=torch.float32)
    with _timed("dwconv3x3_wgrad", B, C, H, W_, xdt):
        L.check(lib.b200ir_dwconv3x3_wgrad(dp, dbs, xp, xbs, xdt, parts.data_ptr(), nparts, B, C, H, W_, _stream()),
                "b200ir_dwconv3x3_wgrad")
    dw, db = reduce_partials_taps(parts, nparts, C, 10, want_bias)
    return dw.view(C, 1, 3, 3), db


def dwconv3x3_gelu_gate_bwd(t, w, bias, dg, want_bias=False, fold=None, sigmoid_gate=False):
    """Backward of the fused dwconv+GELU gate: (da [B,2Hd,H,W] = gradient w.r.t. the depthwise-conv output,
    dw [2Hd,1,3,3], db [2Hd] or None)."""
    B, C2, H, W_ = t.shape
    Hd = C2 // 2
    tp, tbs, tdt = _act(t, "t")
    gp, gbs, gdt = _act(dg, "dg")
    if gdt != tdt:
        raise ValueError("dwconv3x3_gelu_gate_bwd: dtype mismatch")
    da = torch.empty((B, C2, H, W_), device=t.device, dtype=t.dtype)
    ap, abs_, adt = _act(da, "da")
    nparts = lib.b200ir_dwconv3x3_bwd_nparts(B, Hd, H, W_)
    parts = torch.empty((nparts, C2, 10), device=t.device, dtype=torch.float32)
    with _timed("dwconv3x3_gelu_gate_bwd", B, Hd, H, W_, tdt):
        fn = lib.b200ir_dwconv3x3_gelu_sigmoid_gate_bwd if sigmoid_gate else lib.b200ir_dwconv3x3_gelu_gate_bwd
        rc = fn(tp, tbs, _par(w, "w", C2 * 9), _par(bias, "bias", C2), gp, gbs, ap, abs_, tdt, parts.data_ptr(), nparts, B,
                Hd, H, W_, _stream())
    L.check(rc, "b200ir_dwconv3x3_gelu_gate_bwd")
    def finish():
        dw, db = reduce_partials_taps(parts, nparts, C2, 10, want_bias)
        return dw.view(C2, 1, 3, 3), db
    dw, db = fold(finish, parts) if fold is not None else finish()   # see ln_bwd
    return da, dw, db


def mdta_bwd(partials, attn, temperature, w_out, dM):
    """-> (mt [B,C,C], mqk [B,2C,2C], dW_out [C,C], dtemperature [heads]); with a per-sample temperature [B, heads]
    (dynamic temperature) the last item is its gradient [B, heads]."""
    B, heads, nslab, per = partials.shape
    C = w_out.shape[0]
    dev = partials.device
    mt = torch.empty((B, C, C), device=dev, dtype=torch.float32)
    mqk = torch.empty((B, 2 * C, 2 * C), device=dev, dtype=torch.float32)
    dwp = torch.empty((B, C, C), device=dev, dtype=torch.float32)
    dtp = torch.empty((B, heads), device=dev, dtype=torch.float32)
    with _timed("mdta_bwd", B, C, heads):
        per_sample = temperature.dim() == 2
        rc = lib.b200ir_mdta_bwd_ex(partials.data_ptr(), nslab, attn.data_ptr(),
                                    _par(temperature, "temperature", B * heads if per_sample else heads), int(per_sample),
                                    _par(w_out, "w_out", C * C), _par(dM, "dM", B * C * C), mt.data_ptr(), mqk.data_ptr(),
                                    dwp.data_ptr(), dtp.data_ptr(), B, C, heads, _stream())
    L.check(rc, "b200ir_mdta_bwd_ex")
    dw = dwp.view(C, C) if B == 1 else reduce_partials(dwp, B, C * C).view(C, C)
    dt = dtp if per_sample else (dtp.view(heads) if B == 1 else reduce_partials(dtp, B, heads))
    return mt, mqk, dw, dt


def mdta_bwd_chain(partials, attn, temperature, w_out, dM):
    """The softmax / temperature / normalize half of `mdta_bwd` alone: -> (mqk [B,2C,2C], dtemperature)."""
    B, heads, nslab, per = partials.shape
    C = w_out.shape[0]
    mqk = torch.empty((B, 2 * C, 2 * C), device=partials.device, dtype=torch.float32)
    dtp = torch.empty((B, heads), device=partials.device, dtype=torch.float32)
    per_sample = temperature.dim() == 2
    with _timed("mdta_bwd_chain", B, C, heads):
        rc = lib.b200ir_mdta_bwd_parts(partials.data_ptr(), nslab, attn.data_ptr(),
                                       _par(temperature, "temperature", B * heads if per_sample else heads), int(per_sample),
                                       _par(w_out, "w_out", C * C), _par(dM, "dM", B * C * C), None, mqk.data_ptr(), None,
                                       dtp.data_ptr(), B, C, heads, L.MDTA_BWD_CHAIN, _stream())
    L.check(rc, "b200ir_mdta_bwd_parts")
    dt = dtp if per_sample else (dtp.view(heads) if B == 1 else reduce_partials(dtp, B, heads))
    return mqk, dt


def mdta_bwd_fold(attn, w_out, dM, heads):
    """The other half: -> (mt [B,C,C] = the dv operand M_b^T, dW_out [C,C])."""
    C = w_out.shape[0]
    B = dM.shape[0]
    mt = torch.empty((B, C, C), device=dM.device, dtype=torch.float32)
    dwp = torch.empty((B, C, C), device=dM.device, dtype=torch.float32)
    with _timed("mdta_bwd_fold", B, C, heads):
        rc = lib.b200ir_mdta_bwd_parts(None, 1, attn.data_ptr(), None, 0, _par(w_out, "w_out", C * C),
                                       _par(dM, "dM", B * C * C), mt.data_ptr(), None, dwp.data_ptr(), None, B, C, heads,
                                       L.MDTA_BWD_FOLD, _stream())
    L.check(rc, "b200ir_mdta_bwd_parts")
    dw = dwp.view(C, C) if B == 1 else reduce_partials(dwp, B, C * C).view(C, C)
    return mt, dw


def pixel_shuffle(x, mode):
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if mode == L.SHUFFLE2:
        y = torch.empty((B, C // 4, H * 2, W_ * 2), device=x.device, dtype=x.dtype)
    else:
        y = torch.empty((B, C * 4, H // 2, W_ // 2), device=x.device, dtype=x.dtype)
    yp, ybs, ydt = _act(y, "y")
    with _timed("pixel_shuffle", B, C, H, W_, mode, xdt):
        L.check(lib.b200ir_pixel_shuffle(xp, xbs, yp, ybs, xdt, B, C, H, W_, mode, _stream()), "b200ir_pixel_shuffle")
    return y


def l1_loss_fwd(pred, target):
    if pred.shape != target.shape or not (pred.is_contiguous() and target.is_contiguous()):
        raise ValueError("l1_loss: pred/target must be contiguous tensors of the same shape")
    n = pred.numel()
    nparts = min(1024, max(1, (n + 4095) // 4096))
    parts = torch.empty((nparts,), device=pred.device, dtype=torch.float32)
    loss = torch.empty((), device=pred.device, dtype=torch.float32)
    global LAUNCHES
    LAUNCHES += 1
    with _timed("l1_loss_fwd", n):
        rc = lib.b200ir_l1_loss_fwd(pred.data_ptr(), _DT[pred.dtype], target.data_ptr(), _DT[target.dtype], n,
                                    parts.data_ptr(), nparts, loss.data_ptr(), _stream())
    L.check(rc, "b200ir_l1_loss_fwd")
    return loss


def l1_loss_bwd(pred, target, gscale):
    grad = torch.empty_like(pred)
    with _timed("l1_loss_bwd", pred.numel()):
        rc = lib.b200ir_l1_loss_bwd(pred.data_ptr(), _DT[pred.dtype], target.data_ptr(), _DT[target.dtype], pred.numel(),
                                    gscale.data_ptr(), grad.data_ptr(), _DT[grad.dtype], _stream())
    L.check(rc, "b200ir_l1_loss_bwd")
    return grad


def adamw_step(p, g, m, v, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0, step_dev=None, tick=True,
               hyper_dev=None):
    """One AdamW update of a flat buffer (or of a sub-range: pass views).  `step_dev` (int32 CUDA tensor [1]) replaces
    the host `step`: with `tick` the call increments it on the stream and the kernel derives the bias corrections from
    it (CUDA-graph replayable).  `hyper_dev` (float32 CUDA tensor [2] = lr, weight_decay) overrides the host scalars."""
    for t in (p, g, m, v):
        if t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda or t.numel() != p.numel():
            raise ValueError("adamw_step: flat contiguous float32 CUDA buffers of equal size expected")
    if step_dev is not None and (step_dev.dtype != torch.int32 or not step_dev.is_cuda):
        raise ValueError("adamw_step: step_dev must be an int32 CUDA tensor")
    if hyper_dev is not None and (hyper_dev.dtype != torch.float32 or not hyper_dev.is_cuda or hyper_dev.numel() < 2):
        raise ValueError("adamw_step: hyper_dev must be a float32 CUDA tensor [2]")
    global LAUNCHES
    LAUNCHES += 1 if (step_dev is not None and tick) else 0
    with _timed("adamw_step", p.numel()):
        rc = lib.b200ir_adamw_step_ex(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), lr, beta1, beta2,
                                      eps, weight_decay, step, step_dev.data_ptr() if step_dev is not None else None,
                                      int(bool(tick)), hyper_dev.data_ptr() if hyper_dev is not None else None,
                                      grad_scale, _stream())
    L.check(rc, "b200ir_adamw_step")


# ---------------------------------------------------------------------------------------------------
# MoCE-IR decoder pieces (csrc/moce_ops.cu)
# ---------------------------------------------------------------------------------------------------
EW_MUL, EW_SILU_MUL, EW_ADD, EW_GELU, EW_DSILU_MUL = 0, 1, 2, 3, 5


def circconv_patch(a, b, patch, mode=0, out=None):
    """Per (patch x patch) tile and channel: mode 0 circular convolution (= irfft2(rfft2(a) * rfft2(b)), FFTAttention,
    moce_ir.py:406-412), mode 1 circular correlation sum_i a[i] b[i - s] (its gradients)."""
    B, C, H, W_ = a.shape
    if tuple(b.shape) != tuple(a.shape) or b.dtype != a.dtype:
        raise ValueError("circconv_patch: operands must have the same shape and dtype")
    ap, abs_, adt = _act(a, "a")
    bp, bbs, _ = _act(b, "b")
    if out is None:
        out = torch.empty((B, C, H, W_), device=a.device, dtype=a.dtype)
    yp, ybs, _ = _act(out, "out")
    with _timed("circconv_patch", B, C, H, W_, patch, mode, adt):
        rc = lib.b200ir_circconv_patch(ap, abs_, bp, bbs, yp, ybs, adt, B, C, H, W_, patch, mode, _stream())
    L.check(rc, "b200ir_circconv_patch")
    return out


def ew(a, b, op, out=None):
    """Elementwise on NCHW (or [B, N]) views: EW_MUL a*b, EW_SILU_MUL a*silu(b), EW_ADD a+b, EW_GELU gelu(a) (b None),
    EW_DSILU_MUL b * d silu(a)/da."""
    if a.dim() == 2:
        a4 = a.view(a.shape[0], a.shape[1], 1, 1)
        b4 = b.view(b.shape[0], b.shape[1], 1, 1) if b is not None else None
        o4 = out.view(a4.shape) if out is not None else None
        return ew(a4, b4, op, o4).view(a.shape)
    B, C, H, W_ = a.shape
    ap, abs_, adt = _act(a, "a")
    bp, bbs = None, 0
    if b is not None:
        if tuple(b.shape) != tuple(a.shape) or b.dtype != a.dtype:
            raise ValueError("ew: operands must have the same shape and dtype")
        bp, bbs, _ = _act(b, "b")
    if out is None:
        out = torch.empty((B, C, H, W_), device=a.device, dtype=a.dtype)
    yp, ybs, _ = _act(out, "out")
    with _timed("ew", B, C * H * W_, op, adt):
        rc = lib.b200ir_ew(ap, abs_, bp, bbs, yp, ybs, adt, B, C * H * W_, op, _stream())
    L.check(rc, "b200ir_ew")
    return out


def scale_rows(x, g, out=None, accumulate=False):
    """out[b] (+)= g[b] * x[b]; g fp32 [B]."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if out is None:
        if accumulate:
            raise ValueError("scale_rows: accumulate needs an out tensor")
        out = torch.empty((B, C, H, W_), device=x.device, dtype=x.dtype)
    yp, ybs, ydt = _act(out, "out")
    if ydt != xdt:
        raise ValueError("scale_rows: dtype mismatch")
    with _timed("scale_rows", B, C * H * W_, xdt):
        rc = lib.b200ir_scale_rows(xp, xbs, _par(g, "g", B), yp, ybs, xdt, B, C * H * W_, int(accumulate), _stream())
    L.check(rc, "b200ir_scale_rows")
    return out


def spatial_mean(x, gelu=False):
    """fp32 [B, C] = mean over H, W of x (of gelu(x) with gelu=True)."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    out = torch.empty((B, C), device=x.device, dtype=torch.float32)
    with _timed("spatial_mean", B, C, H * W_, xdt):
        rc = lib.b200ir_spatial_mean(xp, xbs, xdt, out.data_ptr(), B, C, H * W_, int(gelu), _stream())
    L.check(rc, "b200ir_spatial_mean")
    return out


def spatial_mean_bwd(g, x, gelu=False):
    """Gradient of spatial_mean w.r.t. x: g fp32 [B, C] -> dx like x."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    dx = torch.empty_like(x, memory_format=torch.contiguous_format)
    dp, dbs, _ = _act(dx, "dx")
    g = g.float().contiguous()
    with _timed("spatial_mean_bwd", B, C, H * W_, xdt):
        rc = lib.b200ir_spatial_mean_bwd(g.data_ptr(), xp, xbs, dp, dbs, xdt, B, C, H * W_, int(gelu), _stream())
    L.check(rc, "b200ir_spatial_mean_bwd")
    return dx


def _rows(t, name):
    """(ptr, row stride, elements per row, dtype code) of a tensor whose leading dimension indexes contiguous rows"""
    if not t.is_cuda or t.dtype not in _DT:
        raise ValueError(f"{name}: expected a float32 / bfloat16 CUDA tensor")
    per = t[0].numel()
    if t.shape[0] > 1 and (t.stride(0) < per or not t[0].is_contiguous()):
        raise ValueError(f"{name}: rows must be contiguous")
    if not t[0].is_contiguous():
        raise ValueError(f"{name}: rows must be contiguous")
    return t.data_ptr(), (t.stride(0) if t.shape[0] > 1 else per), per, _DT[t.dtype]


def _idx(i, n, name):
    if i is None:
        return None
    if i.dtype != torch.int32 or not i.is_cuda or i.numel() != n or not i.is_contiguous():
        raise ValueError(f"{name}: index must be a contiguous int32 CUDA tensor of {n} elements")
    return i.data_ptr()


def rows_copy(x, y, n, src_idx=None, dst_idx=None, scale=None, accumulate=False):
    """y[dst(j)] (+)= scale[j] * x[src(j)] for j < n (rows = samples); SparseDispatcher.dispatch / combine."""
    xp, xrs, per, xdt = _rows(x, "x")
    yp, yrs, pery, ydt = _rows(y, "y")
    if per != pery or xdt != ydt:
        raise ValueError("rows_copy: x / y row size or dtype mismatch")
    sp = None
    if scale is not None:
        if scale.dtype != torch.float32 or scale.numel() != n or not scale.is_contiguous():
            raise ValueError("rows_copy: scale must be contiguous float32 [n]")
        sp = scale.data_ptr()
    with _timed("rows_copy", n, per, xdt):
        rc = lib.b200ir_rows_copy(xp, xrs, yp, yrs, _idx(src_idx, n, "src_idx"), _idx(dst_idx, n, "dst_idx"), sp, n, per,
                                  int(accumulate), xdt, _stream())
    L.check(rc, "b200ir_rows_copy")
    return y


def rowdot(a, b, n, a_idx=None, b_idx=None):
    """fp32 [n]: out[j] = <a[a_idx(j)], b[b_idx(j)]>"""
    ap, ars, per, adt = _rows(a, "a")
    bp, brs, perb, bdt = _rows(b, "b")
    if per != perb or adt != bdt:
        raise ValueError("rowdot: a / b row size or dtype mismatch")
    out = torch.empty((n,), device=a.device, dtype=torch.float32)
    with _timed("rowdot", n, per, adt):
        rc = lib.b200ir_rowdot(ap, ars, _idx(a_idx, n, "a_idx"), bp, brs, _idx(b_idx, n, "b_idx"), out.data_ptr(), n, per,
                               adt, _stream())
    L.check(rc, "b200ir_rowdot")
    return out


# ---------------------------------------------------------------------------------------------------
# AdaIR frequency mining / modulation pieces (csrc/adair_ops.cu)
# ---------------------------------------------------------------------------------------------------
def lowfreq_split(x, hw_counts, max_modes, return_coef=False):
    """FreModule.fft's mask -> ifft2 -> abs (AdaIR-main/net/model.py:335-365): (high, low).  hw_counts int32 [B, 2].
    return_coef: also return the kept spectrum bins (fp32 [B, C, 64, 2]) for `lowfreq_split_bwd`."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    if hw_counts.dtype != torch.int32 or not hw_counts.is_cuda or tuple(hw_counts.shape) != (B, 2):
        raise ValueError("lowfreq_split: hw_counts must be an int32 CUDA tensor [B, 2]")
    hw_counts = hw_counts.contiguous()
    high, low = torch.empty_like(x, memory_format=torch.contiguous_format), torch.empty_like(x, memory_format=torch.contiguous_format)
    hp, hbs, _ = _act(high, "high")
    lp, lbs, _ = _act(low, "low")
    ws = torch.empty((B, C, 64, 2), device=x.device, dtype=torch.float32)
    global LAUNCHES
    LAUNCHES += 1
    with _timed("lowfreq_split", B, C, H, W_, xdt):
        rc = lib.b200ir_lowfreq_split(xp, xbs, hw_counts.data_ptr(), ws.data_ptr(), hp, hbs, lp, lbs, xdt, B, C, H, W_,
                                      max_modes, _stream())
    L.check(rc, "b200ir_lowfreq_split")
    return (high, low, ws) if return_coef else (high, low)


def lowfreq_split_bwd(x, hw_counts, coef, g_high, g_low, max_modes):
    """dx of `lowfreq_split` (csrc/adair_bwd.cu)."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    ghp, ghbs, ghdt = _act(g_high, "g_high")
    glp, glbs, gldt = _act(g_low, "g_low")
    if ghdt != xdt or gldt != xdt or tuple(g_high.shape) != tuple(x.shape) or tuple(g_low.shape) != tuple(x.shape):
        raise ValueError("lowfreq_split_bwd: gradient / input mismatch")
    dx = torch.empty_like(x, memory_format=torch.contiguous_format)
    dp, dbs, _ = _act(dx, "dx")
    ws = torch.empty((B, C, 64, 2), device=x.device, dtype=torch.float32)
    global LAUNCHES
    LAUNCHES += 1
    with _timed("lowfreq_split_bwd", B, C, H, W_, xdt):
        rc = lib.b200ir_lowfreq_split_bwd(xp, xbs, hw_counts.data_ptr(), _par(coef, "coef", B * C * 128), ghp, ghbs, glp, glbs,
                                          ws.data_ptr(), dp, dbs, xdt, B, C, H, W_, max_modes, _stream())
    L.check(rc, "b200ir_lowfreq_split_bwd")
    return dx


def bilinear_resize(x, H, W_):
    B, C, Hi, Wi = x.shape
    xp, xbs, xdt = _act(x, "x")
    y = torch.empty((B, C, H, W_), device=x.device, dtype=x.dtype)
    yp, ybs, _ = _act(y, "y")
    with _timed("bilinear_resize", B, C, Hi, Wi, H, W_, xdt):
        rc = lib.b200ir_bilinear_resize(xp, xbs, yp, ybs, xdt, B, C, Hi, Wi, H, W_, _stream())
    L.check(rc, "b200ir_bilinear_resize")
    return y


def spatial_gate(x, w7, return_mm=False):
    """fp32 [B, H*W] = sigmoid(conv7x7([max_c x, mean_c x])) (SpatialGate); w7 fp32 [1, 2, 7, 7].
    return_mm: also return the [B, 2, H*W] channel max / mean planes for `spatial_gate_bwd`."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    ws = torch.empty((B, 2, H * W_), device=x.device, dtype=torch.float32)
    gate = torch.empty((B, H * W_), device=x.device, dtype=torch.float32)
    global LAUNCHES
    LAUNCHES += 1
    with _timed("spatial_gate", B, C, H, W_, xdt):
        rc = lib.b200ir_spatial_gate(xp, xbs, xdt, _par(w7, "w7", 98), ws.data_ptr(), gate.data_ptr(), B, C, H, W_, _stream())
    L.check(rc, "b200ir_spatial_gate")
    return (gate, ws) if return_mm else gate


def spatial_gate_bwd(x, w7, mm, gate, dgate, dx=None):
    """(dx, dw7 [1, 2, 7, 7]) of `spatial_gate`; a given `dx` is accumulated into."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    acc = dx is not None
    if dx is None:
        dx = torch.empty_like(x, memory_format=torch.contiguous_format)
    dp, dbs, ddt = _act(dx, "dx")
    if ddt != xdt or tuple(dx.shape) != tuple(x.shape):
        raise ValueError("spatial_gate_bwd: dx / x mismatch")
    ds = torch.empty((B, H * W_), device=x.device, dtype=torch.float32)
    dw = torch.empty((1, 2, 7, 7), device=x.device, dtype=torch.float32)
    global LAUNCHES
    LAUNCHES += 2
    with _timed("spatial_gate_bwd", B, C, H, W_, xdt):
        rc = lib.b200ir_spatial_gate_bwd(xp, xbs, xdt, _par(w7, "w7", 98), _par(mm, "mm", B * 2 * H * W_),
                                         _par(gate, "gate", B * H * W_), _par(dgate, "dgate", B * H * W_), ds.data_ptr(), dp, dbs,
                                         dw.data_ptr(), B, C, H, W_, int(acc), _stream())
    L.check(rc, "b200ir_spatial_gate_bwd")
    return dx, dw


def spatial_max(x):
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    out = torch.empty((B, C), device=x.device, dtype=torch.float32)
    with _timed("spatial_max", B, C, H * W_, xdt):
        rc = lib.b200ir_spatial_max(xp, xbs, xdt, out.data_ptr(), B, C, H * W_, _stream())
    L.check(rc, "b200ir_spatial_max")
    return out


def spatial_max_bwd(x, g):
    """dx of `spatial_max`: g [B, C] fp32 at the first maximum of each plane, 0 elsewhere."""
    B, C, H, W_ = x.shape
    xp, xbs, xdt = _act(x, "x")
    dx = torch.empty_like(x, memory_format=torch.contiguous_format)
    dp, dbs, _ = _act(dx, "dx")
    with _timed("spatial_max_bwd", B, C, H * W_, xdt):
        rc = lib.b200ir_spatial_max_bwd(xp, xbs, _par(g, "g", B * C), dp, dbs, xdt, B, C, H * W_, _stream())
    L.check(rc, "b200ir_spatial_max_bwd")
    return dx


def fre_combine(low, high, sw, cw):
    """low * sw[b, p] + high * cw[b, c]; sw fp32 [B, H*W], cw fp32 [B, C]."""
    B, C, H, W_ = low.shape
    lp, lbs, ldt = _act(low, "low")
    hp, hbs, hdt = _act(high, "high")
    if hdt != ldt or tuple(high.shape) != tuple(low.shape):
        raise ValueError("fre_combine: low / high mismatch")
    y = torch.empty((B, C, H, W_), device=low.device, dtype=low.dtype)
    yp, ybs, _ = _act(y, "y")
    with _timed("fre_combine", B, C, H * W_, ldt):
        rc = lib.b200ir_fre_combine(lp, lbs, hp, hbs, _par(sw, "sw", B * H * W_), _par(cw, "cw", B * C), yp, ybs, ldt, B, C,
                                    H * W_, _stream())
    L.check(rc, "b200ir_fre_combine")
    return y


def fre_combine_bwd(g, low, high, sw, cw):
    """(dlow, dhigh, dsw fp32 [B, H*W], dcw fp32 [B, C]) of `fre_combine`."""
    B, C, H, W_ = low.shape
    gp, gbs, gdt = _act(g, "g")
    lp, lbs, ldt = _act(low, "low")
    hp, hbs, hdt = _act(high, "high")
    if not (gdt == ldt == hdt) or tuple(g.shape) != tuple(low.shape) or tuple(high.shape) != tuple(low.shape):
        raise ValueError("fre_combine_bwd: operand mismatch")
    dlow = torch.empty((B, C, H, W_), device=low.device, dtype=low.dtype)
    dhigh = torch.empty((B, C, H, W_), device=low.device, dtype=low.dtype)
    dlp, dlbs, _ = _act(dlow, "dlow")
    dhp, dhbs, _ = _act(dhigh, "dhigh")
    dsw = torch.empty((B, H * W_), device=low.device, dtype=torch.float32)
    dcw = torch.empty((B, C), device=low.device, dtype=torch.float32)
    global LAUNCHES
    LAUNCHES += 1
    with _timed("fre_combine_bwd", B, C, H * W_, ldt):
        rc = lib.b200ir_fre_combine_bwd(gp, gbs, lp, lbs, hp, hbs, _par(sw, "sw", B * H * W_), _par(cw, "cw", B * C), dlp, dlbs,
                                        dhp, dhbs, dsw.data_ptr(), dcw.data_ptr(), ldt, B, C, H * W_, _stream())
    L.check(rc, "b200ir_fre_combine_bwd")
    return dlow, dhigh, dsw, dcw


def affine2(a, b, p1, p2):
    """a * p1[c] + b * p2[c]."""
    B, C, H, W_ = a.shape
    ap, abs_, adt = _act(a, "a")
    bp, bbs, bdt = _act(b, "b")
    if bdt != adt or tuple(b.shape) != tuple(a.shape):
        raise ValueError("affine2: operand mismatch")
    y = torch.empty((B, C, H, W_), device=a.device, dtype=a.dtype)
    yp, ybs, _ = _act(y, "y")
    with _timed("affine2", B, C, H * W_, adt):
        rc = lib.b200ir_affine2(ap, abs_, bp, bbs, _par(p1, "p1", C), _par(p2, "p2", C), yp, ybs, adt, B, C, H * W_, _stream())
    L.check(rc, "b200ir_affine2")
    return y


def affine2_bwd(g, a, b, p1, p2):
    """(da, db, dp1 [C], dp2 [C]) of `affine2`; the per-sample partial sums are folded over B in a fixed order."""
    B, C, H, W_ = a.shape
    gp, gbs, gdt = _act(g, "g")
    ap, abs_, adt = _act(a, "a")
    bp, bbs, bdt = _act(b, "b")
    if not (gdt == adt == bdt) or tuple(g.shape) != tuple(a.shape) or tuple(b.shape) != tuple(a.shape):
        raise ValueError("affine2_bwd: operand mismatch")
    da = torch.empty((B, C, H, W_), device=a.device, dtype=a.dtype)
    db = torch.empty((B, C, H, W_), device=a.device, dtype=a.dtype)
    dap, dabs, _ = _act(da, "da")
    dbp, dbbs, _ = _act(db, "db")
    parts = torch.empty((B, C, 2), device=a.device, dtype=torch.float32)
    with _timed("affine2_bwd", B, C, H * W_, adt):
        rc = lib.b200ir_affine2_bwd(gp, gbs, ap, abs_, bp, bbs, _par(p1, "p1", C), _par(p2, "p2", C), dap, dabs, dbp, dbbs,
                                    parts.data_ptr(), adt, B, C, H * W_, _stream())
    L.check(rc, "b200ir_affine2_bwd")
    dp = reduce_partials(parts.view(B, 2 * C), B, 2 * C).view(C, 2) if B > 1 else parts[0]
    return da, db, dp[:, 0].contiguous(), dp[:, 1].contiguous()
