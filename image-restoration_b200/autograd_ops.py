"""Differentiable wrappers around ``ops`` (one ``torch.autograd.Function`` per fused kernel).

With autograd off (inference, the headline path) these call the forward kernels directly and may
write into caller-provided ``out`` views.  With autograd on, each op records what its hand-written
backward needs (see ``backward.py`` as it lands); until an op's backward exists it raises instead of
silently falling back to PyTorch.
"""
from __future__ import annotations

import torch

from . import _lib as L
from . import ops


def _needs_grad(*ts):
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in ts)


def _no_backward(name):
    raise NotImplementedError(
        f"b200ir: backward of {name} is not implemented yet; run under torch.no_grad() "
        f"(no PyTorch fallback is provided by design)")


def _p(t):
    """fp32 contiguous view of a parameter for the kernels (conv weights keep their memory layout)."""
    if t is None:
        return None
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _sig(*ts):
    return tuple((t.data_ptr(), t._version) if t is not None else None for t in ts)


def packed_weights(cache, key, w, bias, ln_type, ln_w, ln_b):
    """bf16-packed (LayerNorm-folded) copy of a 1x1 conv weight for the tcgen05 GEMM, rebuilt only when a
    parameter changed (optimizer step / load_state_dict bump ``_version``; .to()/.cuda() change the pointer)."""
    sig = _sig(w, bias, ln_w, ln_b) + (ln_type,)
    ent = cache.get(key) if cache is not None else None
    if ent is None or ent[0] != sig:
        cout, cin = w.shape[0], w.shape[1]
        ent = (sig, ops.pack_conv1x1(_p(w).reshape(cout, cin), _p(bias), ln_type, _p(ln_w), _p(ln_b)))
        if cache is not None:
            cache[key] = ent
    return ent[1]


USE_TC = True  # bf16 activations -> tcgen05 GEMMs (set False to force the CUDA-core kernels, e.g. for A/B tests)


def ln_conv1x1(x, w, bias, ln_type, ln_w, ln_b, res, out=None, out_dtype=None, cache=None, key=None):
    if _needs_grad(x, w, bias, ln_w, ln_b, res):
        _no_backward("ln_conv1x1")
    cout = w.shape[0]
    if USE_TC and ops.gemm_tc_usable(x, cout):
        packed = packed_weights(cache, key, w, bias, ln_type, ln_w, ln_b)
        return ops.gemm_tc(x, packed, cout, ln_type, res, out=out, out_dtype=out_dtype)
    return ops.ln_conv1x1(x, _p(w), _p(bias), ln_type, _p(ln_w), _p(ln_b), res, out=out, out_dtype=out_dtype)


def dwconv3x3(x, w, bias):
    if _needs_grad(x, w, bias):
        _no_backward("dwconv3x3")
    return ops.dwconv3x3(x, _p(w), _p(bias))


def dwconv3x3_gelu_gate(x, w, bias):
    if _needs_grad(x, w, bias):
        _no_backward("dwconv3x3_gelu_gate")
    return ops.dwconv3x3_gelu_gate(x, _p(w), _p(bias))


def mdta_core(qkv, temperature, w_out, b_out, heads, res, out=None):
    """Gram+norms -> softmax+fold -> (attn.v + project_out + residual)  (Restormer.py:117-131,147)."""
    if _needs_grad(qkv, temperature, w_out, b_out, res):
        _no_backward("mdta_core")
    B, C3, H, W = qkv.shape
    C = C3 // 3
    partials = ops.mdta_gram(qkv, heads)
    v = qkv[:, 2 * C:]
    if USE_TC and ops.gemm_tc_usable(v, C):
        mp = ops.mdta_softmax_fold(partials, _p(temperature).reshape(-1), _p(w_out).reshape(C, C), packed=True)
        return ops.gemm_tc(v, (mp, _bias_st(b_out, C, mp.shape[1], qkv.device)), C, L.LN_NONE, res, out=out)
    m = ops.mdta_softmax_fold(partials, _p(temperature).reshape(-1), _p(w_out).reshape(C, C))
    return ops.ln_conv1x1(v, m, _p(b_out), L.LN_NONE, None, None, res, out=out, per_sample_w=True)


_ST_CACHE = {}


def _bias_st(bias, C, nrows, device):
    """(s, t) epilogue vectors of the per-sample GEMM: s unused (no LayerNorm), t = project_out bias."""
    key = (None if bias is None else (bias.data_ptr(), bias._version), C, nrows, str(device))
    st = _ST_CACHE.get(key)
    if st is None:
        if len(_ST_CACHE) > 4096:
            _ST_CACHE.clear()
        st = torch.zeros((2, nrows), device=device, dtype=torch.float32)
        if bias is not None:
            st[1, :C] = bias.detach().float()
        _ST_CACHE[key] = st
    return st


def conv3x3(x, w, bias, res, shuffle, out=None, out_dtype=None, cache=None, key=None):
    if _needs_grad(x, w, bias, res):
        _no_backward("conv3x3")
    cout = w.shape[0]
    if USE_TC and res is None and ops.conv3x3_tc_usable(x, cout):
        sig = _sig(w, bias)
        ent = cache.get(key) if cache is not None else None
        if ent is None or ent[0] != sig:
            ent = (sig, ops.pack_conv3x3(_p(w), _p(bias)))
            if cache is not None:
                cache[key] = ent
        return ops.conv3x3_tc(x, ent[1], cout, shuffle, out=out, out_dtype=out_dtype)
    return ops.conv3x3(x, _p(w), _p(bias), res, shuffle, out=out, out_dtype=out_dtype)
