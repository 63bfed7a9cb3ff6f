"""Differentiable wrappers around ``ops`` (one ``torch.autograd.Function`` per fused kernel).

With autograd off (inference, the headline path) these call the forward kernels directly and may
write into caller-provided ``out`` views.  With autograd on, each op goes through a
``torch.autograd.Function`` whose backward launches the hand-written kernels of ``csrc/backward.cu``
(reductions, LayerNorm / gate / softmax backward) and re-uses the forward GEMM / conv kernels with
transposed or flipped weights for the input gradients.  torch's autograd engine only sequences the
calls and sums gradients where a tensor fans out; there is no PyTorch fallback for any op.
"""
from __future__ import annotations

import torch

from . import _lib as L
from . import ops


def _needs_grad(*ts):
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in ts)


def _p(t):
    """fp32 contiguous view of a parameter for the kernels (conv weights keep their memory layout)."""
    if t is None:
        return None
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


PARAM_EPOCH = 0  # bumped by b200ir.AdamW.step(): its kernel updates parameters in place without touching _version


def bump_param_epoch():
    global PARAM_EPOCH
    PARAM_EPOCH += 1


def _optimizer_step_hook(optimizer, args, kwargs):
    """Global torch.optim post-step hook: ANY optimizer step invalidates the packed bf16 weight caches.  Needed because
    a tensor's ``_version`` is not a reliable change signal for every optimizer implementation (measured on B200 with
    torch 2.11: after ``torch.optim.AdamW(fused=True).step()`` the caches keyed on ``(data_ptr, _version)`` served stale
    weights, tests/test_parity_fullsize_gpu.py::test_ddp_wrap_and_stock_adamw)."""
    bump_param_epoch()


try:   # torch >= 2.0
    from torch.optim.optimizer import register_optimizer_step_post_hook as _reg_hook
    _reg_hook(_optimizer_step_hook)
except Exception:   # pragma: no cover  (older torch: rely on _version; invalidate_caches() stays available)
    pass


def invalidate_caches():
    """Drop every packed bf16 weight cache (exported as ``b200ir.invalidate_caches``).  The caches key on
    ``(data_ptr, tensor._version)``; in-place writes through ``p.data`` (EMA ``p.data.mul_().add_()``, re-initialisation
    ``w.data.normal_()``, hand-written optimizers) change neither, so call this after such an update.  Every torch.optim
    optimizer step invalidates the caches by itself (global post-step hook above)."""
    bump_param_epoch()


def _sig(*ts):
    return (PARAM_EPOCH,) + tuple((t.data_ptr(), t._version) if t is not None else None for t in ts)


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
            if _REC is not None and key is not None and w.is_leaf:
                _REC.append((cache, key, w, bias, ln_type, ln_w, ln_b, False))
    return ent[1]


def packed_weights_t(cache, key, w):
    """Operand of the input-gradient GEMM (W^T packed), from the block's cache when a Prepacker filled it this epoch."""
    cout, cin = w.shape[0], w.shape[1]
    if cache is not None and key is not None:
        ent = cache.get(("T", key))
        if ent is not None and ent[0] == _sig(w):
            return ent[1]
    if _REC is not None and cache is not None and key is not None and w.is_leaf:
        _REC.append((cache, key, w, None, L.LN_NONE, None, None, True))
    return ops.pack_conv1x1_t(_p(w).reshape(cout, cin))


# train_step records here every pack that missed its cache during a step ((cache, key, w, bias, ln_type, ln_w, ln_b,
# transposed)); the next steps do all of them in one launch (Prepacker(records=...)), whatever the model is made of
_REC = None


class Prepacker:
    """All 1x1-conv weight packs of a model's Restormer blocks in ONE kernel launch (b200ir_pack_conv1x1_multi): the
    LayerNorm-folded forward operands of qkv / project_in / project_out and the transposed operands of their
    input-gradient GEMMs.  A training step repacks every weight after the optimizer update; done one by one that is
    ~260 launches of 3-5 us.  ``run()`` refreshes the (persistent) packed buffers and stamps the per-module caches with
    the current parameter signature, so that the forward / backward code paths find them instead of packing."""

    def __init__(self, model=None, records=None):
        from . import modules as M
        self.entries = []          # (cache dict, key, (w, bias, ln_w, ln_b), ln_type, packed tuple, transposed)
        self.records = list(records or [])
        descs, blocks = [], 0
        seen = set()

        def add(cache, key, conv, norm, transposed):
            ln_type, lw, lb = (M._ln_args(norm) if (norm is not None and not transposed) else (L.LN_NONE, None, None))
            add_raw(cache, key, conv.weight, None if transposed else conv.bias, ln_type, lw, lb, transposed)

        def add_raw(cache, key, w, bias, ln_type, lw, lb, transposed):
            nonlocal blocks
            tag = (id(cache), key, bool(transposed))
            if tag in seen:
                return
            seen.add(tag)
            cout, cin = w.shape[0], w.shape[1]
            geo = ops.gemm_tc_geometry(cout, cin) if transposed else ops.gemm_tc_geometry(cin, cout)
            if geo is None or w.dtype != torch.float32 or not w.is_cuda:
                return
            kpad, nrows = geo
            wp = torch.empty((1, nrows, kpad), device=w.device, dtype=torch.bfloat16)
            st = torch.empty((2, nrows), device=w.device, dtype=torch.float32)
            d = L.PackDesc()
            d.w = w.data_ptr()
            d.bias = bias.data_ptr() if bias is not None else None
            d.ln_w = lw.data_ptr() if lw is not None else None
            d.ln_b = lb.data_ptr() if (lb is not None and ln_type == L.LN_WITHBIAS) else None
            d.wp, d.st = wp.data_ptr(), st.data_ptr()
            d.Cin, d.Cout, d.Kpad, d.Nrows = cin, cout, kpad, nrows
            d.ln_type, d.transposed, d.first_block, d.reserved = ln_type, int(transposed), blocks, 0
            blocks += ((kpad + 31) // 32) * ((nrows + 31) // 32) if transposed else (nrows + 7) // 8
            descs.append(d)
            self.entries.append((cache, key, (w, bias, lw, lb), ln_type, (wp, st), transposed))

        for rec in self.records:
            add_raw(*rec)
        for m in (model.modules() if model is not None else ()):
            if isinstance(m, (M.TransformerBlock, M.EncoderBlock)):
                if isinstance(m, M.TransformerBlock):
                    n1, attn, n2, ffn = m.norm1, m.attn, m.norm2, m.ffn
                else:
                    n1, attn, n2, ffn = m.norms[0], m.mixer, m.norms[1], m.ffn
                ca = attn.__dict__.setdefault("_b200ir_cache", {})
                cf = ffn.__dict__.setdefault("_b200ir_cache", {})
                for cache, key, conv, norm in ((ca, ("qkv", id(n1)), attn.qkv, n1), (cf, ("in", id(n2)), ffn.project_in, n2),
                                               (cf, "out", ffn.project_out, None)):
                    add(cache, key, conv, norm, False)
                    add(cache, key, conv, None, True)
        self.ndesc, self.total_blocks = len(descs), blocks
        self.descs_dev = None
        if descs:
            arr = (L.PackDesc * len(descs))(*descs)
            host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            self.descs_dev = host.to(self.entries[0][2][0].device)
        self.ptrs = self._pointers()

    def _pointers(self):
        return tuple(t.data_ptr() if t is not None else 0 for e in self.entries for t in e[2])

    def valid(self):
        """EVERY parameter the table captured (weights, biases, LayerNorm weight / bias) still lives where the
        descriptors say (no .to()/.cuda()/re-flattening since construction)."""
        return self.descs_dev is not None and self.ptrs == self._pointers()

    def run(self):
        if not self.valid():
            return False
        ops.pack_conv1x1_multi(self.descs_dev, self.ndesc, self.total_blocks)
        for cache, key, (w, bias, lw, lb), ln_type, packed, transposed in self.entries:
            if transposed:
                cache[("T", key)] = (_sig(w), packed)
            else:
                cache[key] = (_sig(w, bias, lw, lb) + (ln_type,), packed)
        return True


USE_TC = True  # bf16 activations -> tcgen05 GEMMs (set False to force the CUDA-core kernels, e.g. for A/B tests)


def ln_conv1x1(x, w, bias, ln_type, ln_w, ln_b, res, out=None, out_dtype=None, cache=None, key=None):
    if _needs_grad(x, w, bias, ln_w, ln_b, res):
        return _LnConv1x1.apply(x, w, bias, ln_w, ln_b, res, ln_type, out_dtype, cache, key)
    return _ln_conv1x1_fwd(x, w, bias, ln_type, ln_w, ln_b, res, out, out_dtype, cache, key)


def _ln_conv1x1_fwd(x, w, bias, ln_type, ln_w, ln_b, res, out=None, out_dtype=None, cache=None, key=None):
    cout = w.shape[0]
    if USE_TC and ops.gemm_tc_usable(x, cout):
        packed = packed_weights(cache, key, w, bias, ln_type, ln_w, ln_b)
        return ops.gemm_tc(x, packed, cout, ln_type, res, out=out, out_dtype=out_dtype)
    return ops.ln_conv1x1(x, _p(w), _p(bias), ln_type, _p(ln_w), _p(ln_b), res, out=out, out_dtype=out_dtype)


USE_FUSED = True  # inference: LN + 1x1 + depthwise 3x3 (+gate) as one kernel where the shape is supported (fused_ln_proj_dw.cu)


def fused_ln_proj_dw(x, conv_w, conv_b, dw_w, dw_b, ln_type, ln_w, ln_b, cout, gate, cache=None, key=None):
    """out = [gate of] dwconv3x3(conv1x1(LN(x))) in ONE kernel, or None when the fused kernel does not apply (autograd
    needed, fp32 activations, unsupported shape): the caller then runs the two-kernel path.  Reference:
    Restormer.py:88-91 (gate = True) / :114 (gate = False)."""
    if not USE_FUSED or ln_type == L.LN_NONE or _needs_grad(x, conv_w, conv_b, dw_w, dw_b, ln_w, ln_b):
        return None
    if not ops.ln_proj_dw_usable(x, cout, gate):
        return None
    sig = _sig(conv_w, conv_b, ln_w, ln_b) + (ln_type, "fused")
    ent = cache.get(key) if cache is not None else None
    if ent is None or ent[0] != sig:
        ctot, cin = conv_w.shape[0], conv_w.shape[1]
        ent = (sig, ops.pack_fused_proj_dw(_p(conv_w).reshape(ctot, cin), _p(conv_b), ln_type, _p(ln_w), _p(ln_b), cout, gate))
        if cache is not None:
            cache[key] = ent
    return ops.ln_proj_dw(x, ent[1], _p(dw_w), _p(dw_b), ln_type, cout, gate)


def dwconv3x3(x, w, bias):
    if _needs_grad(x, w, bias):
        return _DwConv.apply(x, w, bias)
    return ops.dwconv3x3(x, _p(w), _p(bias))


def dwconv3x3_gelu_gate(x, w, bias, sigmoid_gate=False):
    if _needs_grad(x, w, bias):
        return _DwGate.apply(x, w, bias, sigmoid_gate)
    if sigmoid_gate:
        return ops.dwconv3x3_gelu_sigmoid_gate(x, _p(w), _p(bias))
    return ops.dwconv3x3_gelu_gate(x, _p(w), _p(bias))


def mdta_core(qkv, temperature, w_out, b_out, heads, res, out=None, cache=None, dynamic_temp=None):
    """Gram+norms -> softmax+fold -> (attn.v + project_out + residual)  (Restormer.py:117-131,147).
    `dynamic_temp` ([B] or [B,1,1,1]): per-sample factor on the temperature
    (experiment/RectiFormer.py:309-311)."""
    if dynamic_temp is not None:
        if _needs_grad(qkv, temperature, w_out, b_out, res, dynamic_temp):
            # effective temperature [B, heads] = temperature[h] * dynamic_temp[b]: a torch op on B * heads numbers, so that
            # autograd splits the kernel's d(effective temperature) between the parameter and the dynamic factor
            eff = temperature.float().reshape(1, -1) * dynamic_temp.float().reshape(-1, 1)
            return _MdtaCore.apply(qkv, eff, w_out, b_out, res, heads, cache)
        temperature = (_p(temperature).reshape(1, -1) * dynamic_temp.detach().float().reshape(-1, 1)).contiguous()
        return _mdta_core_fwd(qkv, temperature, w_out, b_out, heads, res, out, cache)[0]
    if _needs_grad(qkv, temperature, w_out, b_out, res):
        return _MdtaCore.apply(qkv, temperature, w_out, b_out, res, heads, cache)
    return _mdta_core_fwd(qkv, temperature, w_out, b_out, heads, res, out, cache)[0]


def _per_sample_gemm(x, m, bias, res, out=None):
    """y = res + bias + m[b] . x[b] with a per-sample fp32 weight m [B,Cout,Cin] (tcgen05 when bf16)."""
    cout = m.shape[1]
    if USE_TC and ops.gemm_tc_usable(x, cout):
        wp, st = ops.pack_conv1x1(m, bias, per_sample=True)
        return ops.gemm_tc(x, (wp, st), cout, L.LN_NONE, res, out=out)
    return ops.ln_conv1x1(x, m, bias, L.LN_NONE, None, None, res, out=out, per_sample_w=True)


def _mdta_core_fwd(qkv, temperature, w_out, b_out, heads, res, out=None, cache=None):
    B, C3, H, W = qkv.shape
    C = C3 // 3
    partials = ops.mdta_gram(qkv, heads)
    v = qkv[:, 2 * C:]
    if USE_TC and ops.gemm_tc_usable(v, C):
        mp, attn = ops.mdta_softmax_fold(partials, _temp(temperature), _p(w_out).reshape(C, C), packed=True, want_attn=True)
        y = ops.gemm_tc(v, (mp, _bias_st(b_out, C, mp.shape[1], qkv.device, cache)), C, L.LN_NONE, res, out=out)
    else:
        m, attn = ops.mdta_softmax_fold(partials, _temp(temperature), _p(w_out).reshape(C, C), want_attn=True)
        y = ops.ln_conv1x1(v, m, _p(b_out), L.LN_NONE, None, None, res, out=out, per_sample_w=True)
    return y, partials, attn


def _temp(temperature):
    """[heads] parameter view, or the [B, heads] per-sample (dynamic) temperature as it is."""
    t = _p(temperature)
    return t if t.dim() == 2 else t.reshape(-1)


def _bias_st(bias, C, nrows, device, cache=None):
    """(s, t) epilogue vectors of the per-sample GEMM: s unused (no LayerNorm), t = project_out bias.  Cached only in
    the owning module's cache (a tensor's address alone does not identify it once it has been freed)."""
    # (without a bias the vectors are zeros whatever the parameters are: no parameter epoch in the key, or every optimizer
    # step would refill them -- 44 fill launches per training step)
    sig = (_sig(bias) if bias is not None else ("nobias",)) + (C, nrows)
    ent = cache.get("bias_st") if cache is not None else None
    if ent is None or ent[0] != sig:
        st = torch.zeros((2, nrows), device=device, dtype=torch.float32)
        if bias is not None:
            st[1, :C] = bias.detach().float()
        ent = (sig, st)
        if cache is not None:
            cache["bias_st"] = ent
    return ent[1]


def conv3x3(x, w, bias, res, shuffle, out=None, out_dtype=None, cache=None, key=None):
    if _needs_grad(x, w, bias, res):
        return _Conv3x3.apply(x, w, bias, res, shuffle, out_dtype, cache, key)
    return _conv3x3_fwd(x, w, bias, res, shuffle, out, out_dtype, cache, key)


def _conv3x3_fwd(x, w, bias, res, shuffle, out=None, out_dtype=None, cache=None, key=None):
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


# ---------------------------------------------------------------------------------------------------
# autograd.Functions: forward = the kernels above, backward = csrc/backward.cu + forward kernels re-used
# ---------------------------------------------------------------------------------------------------
def _cg(g):
    """Incoming gradient as an NCHW view our kernels accept (autograd may hand over expanded / permuted views)."""
    s = g.stride()
    b, c, h, w = g.shape
    ok = (s[3] == 1 or w == 1) and (s[2] == w or h == 1) and (s[1] == h * w or c == 1)
    return g if ok else g.contiguous()


# ---- weight gradients on a side stream -------------------------------------------------------------------------------
# Weight / bias gradients are leaves of the backward pass: nothing downstream of a layer's backward needs them before the
# optimizer.  Inside b200ir.train_step / GraphedTrainStep (SIDE_WGRAD set there and nowhere else: a generic
# loss.backward() under torch DDP reads gradients from hooks that know nothing about this stream) they are enqueued on a
# second stream forked from the main one, so that the many small launches (split-K folds, deep-level weight-gradient GEMMs
# with a handful of tiles) fill SMs the input-gradient chain leaves idle; captured in a CUDA graph this becomes a parallel
# branch.  The arithmetic and its order are unchanged: results are bitwise those of the single-stream path.
SIDE_WGRAD = False
_SIDE_STREAMS = {}
_SIDE_USED = False
# References to the tensors the side stream reads, held until join_side(): autograd's gradient accumulation adds IN PLACE
# into a buffered gradient nobody else references (input_buffer.cpp can_accumulate_inplace), e.g. into the dy that a
# residual layer hands on as the gradient of its skip input -- on the main stream, while the side stream still reads it.
_SIDE_KEEP = []
_SIDE_GRADS = []   # (parameter, gradient computed on the side stream), assigned at join_side()
SIDE_SITES = {"conv1x1", "dw3x3", "conv3x3", "lnfold", "gatefold"}   # which weight gradients fork (A/B, debugging)
FORK_ATTN_BWD = True   # the two independent halves of the attention backward on two streams (inside train_step only)
_AUX_STREAMS = {}


def _aux_stream(dev):
    s = _AUX_STREAMS.get(dev)
    if s is None:
        s = _AUX_STREAMS[dev] = torch.cuda.Stream(device=dev)
    return s


class _side:
    """``with _side(site, (w, bias), in1, in2) as sd: dw = sd.give(w, f(in1, in2))`` runs f's launches on the side stream
    after everything enqueued so far on the current stream; inputs are kept alive and marked for the allocator until
    join_side(), which also assigns the gradients to their (leaf) parameters."""

    def __init__(self, site, params, *inputs):
        self.site, self.params, self.inputs, self.cm = site, params, inputs, None

    def __enter__(self):
        global _SIDE_USED
        if not SIDE_WGRAD or self.site not in SIDE_SITES:
            return self
        if any(q is not None and not (q.is_leaf and q.requires_grad) for q in self.params):
            return self      # a computed (non-leaf) weight: its gradient has to flow on through autograd
        self.main = torch.cuda.current_stream()
        dev = self.main.device.index
        self.side = _SIDE_STREAMS.get(dev)
        if self.side is None:
            self.side = _SIDE_STREAMS[dev] = torch.cuda.Stream(device=dev)
        self.side.wait_stream(self.main)
        for t in self.inputs:
            if t is not None:
                t.record_stream(self.side)
                _SIDE_KEEP.append(t)
        self.cm = torch.cuda.stream(self.side)
        self.cm.__enter__()
        _SIDE_USED = True
        return self

    @property
    def active(self):
        return self.cm is not None

    def give(self, param, grad):
        """hand a gradient computed on the side stream to its parameter at join_side() instead of returning it to
        autograd (AccumulateGrad may clone or add it on the MAIN stream as soon as backward() returns); -> None"""
        if grad is None or param is None:
            return None
        if self.cm is None:
            return grad
        grad.record_stream(self.main)
        _SIDE_GRADS.append((param, grad))
        return None

    def __exit__(self, *exc):
        if self.cm is not None:
            self.cm.__exit__(*exc)
        return False


def _side_fold(site, params):
    """`fold` argument for ops.ln_bwd / ops.dwconv3x3_gelu_gate_bwd: run the fold of the parameter-gradient partial sums on
    the side stream.  -> (callable, holder); holder.sd is the _side context the gradients are given through."""
    class _H:
        sd = None
    h = _H()

    def fold(fn, parts):
        h.sd = _side(site, params, parts)
        with h.sd:
            return fn()
    return fold, h


def join_side():
    """make the current stream wait for the side-stream work enqueued since the last join"""
    global _SIDE_USED
    if _SIDE_USED:
        main = torch.cuda.current_stream()
        side = _SIDE_STREAMS.get(main.device.index)
        if side is not None:
            main.wait_stream(side)
        _SIDE_USED = False
    _SIDE_KEEP.clear()
    for prm, g in _SIDE_GRADS:
        g = g.view(prm.shape) if g.shape != prm.shape else g
        if g.dtype != prm.dtype:
            g = g.to(prm.dtype)
        prm.grad = g if prm.grad is None else prm.grad + g
    _SIDE_GRADS.clear()


class _LnConv1x1(torch.autograd.Function):
    """y = res + bias + W.LN(x)  (Restormer.py:37-39,54-57 + the 1x1 nn.Conv2d + :147-148)."""

    @staticmethod
    def forward(ctx, x, w, bias, ln_w, ln_b, res, ln_type, out_dtype, cache, key):
        y = _ln_conv1x1_fwd(x, w, bias, ln_type, ln_w, ln_b, res, None, out_dtype, cache, key)
        ctx.save_for_backward(x, w, ln_w, ln_b)
        ctx.ln_type, ctx.has_bias, ctx.has_res = ln_type, bias is not None, res is not None
        ctx.cache, ctx.key = cache, key
        ctx.params, ctx.lnparams = (w, bias), (ln_w, ln_b)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, ln_w, ln_b = ctx.saved_tensors
        dy = _cg(dy)
        cout, cin = w.shape[0], w.shape[1]
        need = ctx.needs_input_grad
        dx = dlw = dlb = None
        xn = x if ctx.ln_type == L.LN_NONE else None
        if need[0] or ctx.ln_type != L.LN_NONE:
            if USE_TC and ops.gemm_tc_usable(dy, cin) and x.dtype == torch.bfloat16:
                dn = ops.gemm_tc(dy, packed_weights_t(ctx.cache, ctx.key, w), cin, L.LN_NONE, None, out_dtype=x.dtype)
            else:
                wt = _p(w).reshape(cout, cin).t().contiguous()
                dn = ops.ln_conv1x1(dy, wt, None, L.LN_NONE, None, None, None, out_dtype=x.dtype)
            if ctx.ln_type != L.LN_NONE:
                # one pass over (x, dn): dx, dgamma/dbeta and LN(x) itself (operand of the weight gradient below)
                fold, h = _side_fold("lnfold", ctx.lnparams)
                dx, dlw, dlb, xn = ops.ln_bwd(x, dn, ctx.ln_type, _p(ln_w), None, _p(ln_b), want_xn=True, fold=fold)
                if ctx.ln_type != L.LN_WITHBIAS:
                    dlb = None
                dlw, dlb = h.sd.give(ctx.lnparams[0], dlw), h.sd.give(ctx.lnparams[1], dlb)
            else:
                dx = dn
        with _side("conv1x1", ctx.params, dy, xn) as sd:
            dw, db = ops.conv_wgrad(dy, xn, 1, want_bias=ctx.has_bias)
            dw, db = sd.give(ctx.params[0], dw.view(w.shape).to(w.dtype)), sd.give(ctx.params[1], db)
        del xn
        dres = dy if ctx.has_res else None
        return (dx, dw, db, dlw, dlb, dres, None, None, None, None)


class _LayerNorm(torch.autograd.Function):
    """Stand-alone channel LayerNorm (Restormer.py:60-70 called outside a TransformerBlock, e.g. the 18 norms
    dropin.accelerate() adopts inside MoCE-IR's decoder)."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, ln_type):
        ctx.save_for_backward(x, ln_w, ln_b)
        ctx.ln_type = ln_type
        return ops.ln_fwd(x, ln_type, _p(ln_w), _p(ln_b))

    @staticmethod
    def backward(ctx, dy):
        x, ln_w, ln_b = ctx.saved_tensors
        dx, dlw, dlb = ops.ln_bwd(x, _cg(dy).to(x.dtype), ctx.ln_type, _p(ln_w), None, _p(ln_b))
        return dx, dlw.to(ln_w.dtype), (dlb.to(ln_b.dtype) if ln_b is not None else None), None


def layernorm(x, ln_type, ln_w, ln_b):
    x = _cg(x)
    if _needs_grad(x, ln_w, ln_b):
        return _LayerNorm.apply(x, ln_w, ln_b, ln_type)
    return ops.ln_fwd(x, ln_type, _p(ln_w), _p(ln_b))


class _DwConv(torch.autograd.Function):
    """Depthwise 3x3 (Restormer.py:106,114)."""

    @staticmethod
    def forward(ctx, x, w, bias):
        ctx.save_for_backward(x, w)
        ctx.has_bias = bias is not None
        ctx.params = (w, bias)
        return ops.dwconv3x3(x, _p(w), _p(bias))

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = _cg(dy)
        with _side("dw3x3", ctx.params, dy, x) as sd:
            dw, db = ops.dwconv3x3_wgrad(dy, x, want_bias=ctx.has_bias)
            dw, db = sd.give(ctx.params[0], dw.to(w.dtype)), sd.give(ctx.params[1], db)
        dx = ops.dwconv3x3_bwd_input(dy, _p(w)) if ctx.needs_input_grad[0] else None
        return dx, dw, db


class _DwConvKxK(torch.autograd.Function):
    """Depthwise K x K (K = 3, 5, 7), e.g. CrossAttention.kv_dwconv (moce_ir.py:341)."""

    @staticmethod
    def forward(ctx, x, w, bias):
        ctx.save_for_backward(x, w)
        ctx.has_bias = bias is not None
        return ops.dwconv_kxk(x, _p(w), _p(bias))

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dx, dw, db = ops.dwconv_kxk_bwd(_cg(dy), x, _p(w), want_bias=ctx.has_bias, want_input=ctx.needs_input_grad[0])
        return dx, dw.to(w.dtype), db


def dwconv_kxk(x, w, bias):
    if _needs_grad(x, w, bias):
        return _DwConvKxK.apply(x, w, bias)
    return ops.dwconv_kxk(x, _p(w), _p(bias))


class _SigmoidGate(torch.autograd.Function):
    """y = u * sigmoid(g) with a head-wise or element-wise gate (gated MDTA variants)."""

    @staticmethod
    def forward(ctx, u, g):
        ctx.save_for_backward(u, g)
        return ops.sigmoid_gate(u, g)

    @staticmethod
    def backward(ctx, dy):
        u, g = ctx.saved_tensors
        return ops.sigmoid_gate_bwd(u, g, _cg(dy))


def sigmoid_gate(u, g):
    if _needs_grad(u, g):
        return _SigmoidGate.apply(u, g)
    return ops.sigmoid_gate(u, g)


class _DwGate(torch.autograd.Function):
    """dwconv + chunk + gelu(x1)*x2 (Restormer.py:84,90-91); the conv output is recomputed in backward."""

    @staticmethod
    def forward(ctx, t, w, bias, sigmoid_gate=False):
        ctx.save_for_backward(t, w, bias)
        ctx.params = (w, bias)
        ctx.sigmoid_gate = sigmoid_gate   # gelu(x1) * sigmoid(x2): the split-gate FFN of experiment/RectiFormer.py:231-266
        if sigmoid_gate:
            return ops.dwconv3x3_gelu_sigmoid_gate(t, _p(w), _p(bias))
        return ops.dwconv3x3_gelu_gate(t, _p(w), _p(bias))

    @staticmethod
    def backward(ctx, dg):
        t, w, bias = ctx.saved_tensors
        dg = _cg(dg)
        fold, h = _side_fold("gatefold", ctx.params)
        da, dw, db = ops.dwconv3x3_gelu_gate_bwd(t, _p(w), _p(bias), dg, want_bias=bias is not None, fold=fold,
                                                 sigmoid_gate=ctx.sigmoid_gate)
        dw, db = h.sd.give(ctx.params[0], dw.to(w.dtype)), h.sd.give(ctx.params[1], db)
        dt = ops.dwconv3x3_bwd_input(da, _p(w)) if ctx.needs_input_grad[0] else None
        return dt, dw, db, None


class _MdtaCore(torch.autograd.Function):
    """F.normalize + q.k^T*temperature + softmax + attn.v + project_out + residual (Restormer.py:117-131,147)."""

    @staticmethod
    def forward(ctx, qkv, temperature, w_out, b_out, res, heads, cache=None):
        y, partials, attn = _mdta_core_fwd(qkv, temperature, w_out, b_out, heads, res, None, cache)
        ctx.save_for_backward(qkv, temperature, w_out, partials, attn)
        ctx.has_bias, ctx.has_res, ctx.heads = b_out is not None, res is not None, heads
        return y

    @staticmethod
    def backward(ctx, dout):
        qkv, temperature, w_out, partials, attn = ctx.saved_tensors
        dout = _cg(dout)
        B, C3, H, W = qkv.shape
        C = C3 // 3
        v = qkv[:, 2 * C:]
        dM, db = ops.conv_wgrad(dout, v, 1, per_sample=True, want_bias=ctx.has_bias)
        dM = dM.contiguous()
        dqkv = torch.empty_like(qkv)
        if SIDE_WGRAD and FORK_ATTN_BWD:
            # inside b200ir.train_step: the two halves only share inputs -- {M_b^T, dW_out, dv = M_b^T . dout} on a second
            # stream, {softmax / temperature / normalize chain, d[q;k] = mqk_b . [q;k]} on this one; joined before returning
            main = torch.cuda.current_stream()
            aux = _aux_stream(main.device.index)
            aux.wait_stream(main)
            with torch.cuda.stream(aux):
                mt, dw = ops.mdta_bwd_fold(attn, _p(w_out).reshape(C, C), dM, ctx.heads)
                _per_sample_gemm(dout, mt, None, None, out=dqkv[:, 2 * C:])
                dw.record_stream(main)
                del mt
            mqk, dtemp = ops.mdta_bwd_chain(partials, attn, _temp(temperature), _p(w_out).reshape(C, C), dM)
            _per_sample_gemm(qkv[:, :2 * C], mqk, None, None, out=dqkv[:, :2 * C])
            main.wait_stream(aux)
        else:
            mt, mqk, dw, dtemp = ops.mdta_bwd(partials, attn, _temp(temperature), _p(w_out).reshape(C, C), dM)
            _per_sample_gemm(dout, mt, None, None, out=dqkv[:, 2 * C:])
            _per_sample_gemm(qkv[:, :2 * C], mqk, None, None, out=dqkv[:, :2 * C])
        return (dqkv, dtemp.view(temperature.shape).to(temperature.dtype), dw.view(w_out.shape).to(w_out.dtype), db,
                dout if ctx.has_res else None, None, None)


class _Conv3x3(torch.autograd.Function):
    """Dense 3x3 conv (+bias, +res) with PixelUnshuffle/PixelShuffle folded (Restormer.py:162-189,281)."""

    @staticmethod
    def forward(ctx, x, w, bias, res, shuffle, out_dtype, cache, key):
        y = _conv3x3_fwd(x, w, bias, res, shuffle, None, out_dtype, cache, key)
        ctx.save_for_backward(x, w)
        ctx.params = (w, bias)
        ctx.has_bias, ctx.has_res, ctx.shuffle = bias is not None, res is not None, shuffle
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = _cg(dy)
        need = ctx.needs_input_grad
        if ctx.shuffle == L.UNSHUFFLE2:      # y = unshuffle(conv(x)) -> dconv = shuffle(dy)
            dc = ops.pixel_shuffle(dy, L.SHUFFLE2)
        elif ctx.shuffle == L.SHUFFLE2:
            dc = ops.pixel_shuffle(dy, L.UNSHUFFLE2)
        else:
            dc = dy
        with _side("conv3x3", ctx.params, dc, x) as sd:
            dw, db = ops.conv_wgrad(dc, x, 9, want_bias=ctx.has_bias)
            dw, db = sd.give(ctx.params[0], dw.view(w.shape).to(w.dtype)), sd.give(ctx.params[1], db)
        dx = None
        if need[0]:
            cin = w.shape[1]
            wt = _p(w).transpose(0, 1).flip(2, 3).contiguous()
            if USE_TC and ops.conv3x3_tc_usable(dc, cin) and x.dtype == torch.bfloat16:
                dx = ops.conv3x3_tc(dc, ops.pack_conv3x3(wt, None), cin, L.SHUFFLE_NONE, out_dtype=x.dtype)
            else:
                dx = ops.conv3x3(dc, wt, None, None, L.SHUFFLE_NONE, out_dtype=x.dtype)
        dres = dy if (ctx.has_res and need[3]) else None
        return dx, dw, db, dres, None, None, None, None


class _L1Loss(torch.autograd.Function):
    """nn.L1Loss() (mean reduction), MoCE-IR-main/src/train.py:51,65."""

    @staticmethod
    def forward(ctx, pred, target):
        pred, target = pred.contiguous(), target.contiguous()
        ctx.save_for_backward(pred, target)
        return ops.l1_loss_fwd(pred, target)

    @staticmethod
    def backward(ctx, g):
        pred, target = ctx.saved_tensors
        return ops.l1_loss_bwd(pred, target, g.contiguous().float()), None


def l1_loss(pred, target):
    if _needs_grad(pred):
        return _L1Loss.apply(pred, target)
    return ops.l1_loss_fwd(pred.contiguous(), target.contiguous())


# ---- MoCE-IR adapter pieces (moce_ir.py:71-141, 373-422, 548-560, 659-667, 703-707, 1068-1075) --------------------
class _CircConv(torch.autograd.Function):
    """Per-patch circular convolution irfft2(rfft2(a) * rfft2(b)) (FFTAttention, moce_ir.py:406-412); the gradients are
    the circular correlations of the output gradient with the other operand."""

    @staticmethod
    def forward(ctx, a, b, patch):
        ctx.save_for_backward(a, b)
        ctx.patch = patch
        return ops.circconv_patch(a, b, patch)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        g = _cg(g).to(a.dtype)
        da = ops.circconv_patch(g, b, ctx.patch, mode=1) if ctx.needs_input_grad[0] else None
        db = ops.circconv_patch(g, a, ctx.patch, mode=1) if ctx.needs_input_grad[1] else None
        return da, db, None


def circconv_patch(a, b, patch):
    if _needs_grad(a, b):
        return _CircConv.apply(a, b, patch)
    return ops.circconv_patch(a, b, patch)


class _EwMul(torch.autograd.Function):
    """y = a * b (FFTAttention: LayerNorm(out) * v, moce_ir.py:418)."""

    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return ops.ew(a, b, ops.EW_MUL)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        g = _cg(g).to(a.dtype)
        return (ops.ew(g, b, ops.EW_MUL) if ctx.needs_input_grad[0] else None,
                ops.ew(g, a, ops.EW_MUL) if ctx.needs_input_grad[1] else None)


class _EwSiluMul(torch.autograd.Function):
    """y = a * silu(b) (ModExpert.process, moce_ir.py:556-558)."""

    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return ops.ew(a, b, ops.EW_SILU_MUL)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        g = _cg(g).to(a.dtype)
        da = ops.ew(g, b, ops.EW_SILU_MUL) if ctx.needs_input_grad[0] else None
        db = ops.ew(b, ops.ew(g, a, ops.EW_MUL), ops.EW_DSILU_MUL) if ctx.needs_input_grad[1] else None
        return da, db


def ew_mul(a, b):
    return _EwMul.apply(a, b) if _needs_grad(a, b) else ops.ew(a, b, ops.EW_MUL)


def ew_silu_mul(a, b):
    return _EwSiluMul.apply(a, b) if _needs_grad(a, b) else ops.ew(a, b, ops.EW_SILU_MUL)


class _SpatialMean(torch.autograd.Function):
    """fp32 [B, C] = mean over (H, W) of x, or of gelu(x): AdaptiveAvgPool2d(1) of the router (moce_ir.py:703-707) and
    FrequencyEmbedding's gelu(...).mean(dim=(-2, -1)) (:1068-1073)."""

    @staticmethod
    def forward(ctx, x, gelu):
        ctx.save_for_backward(x)
        ctx.gelu = gelu
        return ops.spatial_mean(x, gelu=gelu)

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        return ops.spatial_mean_bwd(g, x, gelu=ctx.gelu), None


def spatial_mean(x, gelu=False):
    x = _cg(x)
    return _SpatialMean.apply(x, gelu) if _needs_grad(x) else ops.spatial_mean(x, gelu=gelu)


# ---------------------------------------------------------------------------------------------------------------------
# AdaIR frequency mining / modulation under autograd (csrc/adair_bwd.cu; AdaIR-main/net/model.py:234-372)
# ---------------------------------------------------------------------------------------------------------------------
class _SpatialMax(torch.autograd.Function):
    """AdaptiveMaxPool2d(1) of ChannelGate (model.py:253): the gradient goes to the first maximum of each plane."""

    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return ops.spatial_max(x)

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        return ops.spatial_max_bwd(x, g.float().contiguous())


def spatial_max(x):
    x = _cg(x)
    return _SpatialMax.apply(x) if _needs_grad(x) else ops.spatial_max(x)


class _LowfreqSplit(torch.autograd.Function):
    """FreModule.fft's fft2 -> mask -> ifft2 -> abs (model.py:340-372): (high, low).  hw_counts carries no gradient (the
    thresholds pass through .int() in the reference)."""

    @staticmethod
    def forward(ctx, x, hw_counts, max_modes):
        high, low, coef = ops.lowfreq_split(x, hw_counts, max_modes, return_coef=True)
        ctx.save_for_backward(x, hw_counts, coef)
        ctx.max_modes = max_modes
        return high, low

    @staticmethod
    def backward(ctx, g_high, g_low):
        x, hw_counts, coef = ctx.saved_tensors
        return ops.lowfreq_split_bwd(x, hw_counts, coef, _cg(g_high), _cg(g_low), ctx.max_modes), None, None


def lowfreq_split(x, hw_counts, max_modes):
    x = _cg(x)
    return _LowfreqSplit.apply(x, hw_counts, max_modes) if _needs_grad(x) else ops.lowfreq_split(x, hw_counts, max_modes)


class _FreRefineGate(torch.autograd.Function):
    """FreRefine up to its projection (model.py:279-285): low * SpatialGate(high) + high * channel_weight, with the 7x7
    SpatialGate convolution inside (weight w7); channel_weight cw fp32 [B, C] comes from ChannelGate's tiny MLP."""

    @staticmethod
    def forward(ctx, low, high, w7, cw):
        w = _p(w7)
        gate, mm = ops.spatial_gate(high, w, return_mm=True)
        cwc = cw.detach().float().contiguous()
        ctx.save_for_backward(low, high, w, gate, mm, cwc)
        return ops.fre_combine(low, high, gate, cwc)

    @staticmethod
    def backward(ctx, g):
        low, high, w, gate, mm, cw = ctx.saved_tensors
        dlow, dhigh, dsw, dcw = ops.fre_combine_bwd(_cg(g), low, high, gate, cw)
        dhigh, dw7 = ops.spatial_gate_bwd(high, w, mm, gate, dsw, dx=dhigh)
        return dlow, dhigh, dw7, dcw


def fre_refine_gate(low, high, w7, cw):
    low, high = _cg(low), _cg(high)
    if _needs_grad(low, high, w7, cw):
        return _FreRefineGate.apply(low, high, w7, cw)
    return ops.fre_combine(low, high, ops.spatial_gate(high, _p(w7)), cw.float().contiguous())


class _Affine2(torch.autograd.Function):
    """out * para1 + y * para2 (model.py:324), para [C, 1, 1]."""

    @staticmethod
    def forward(ctx, a, b, p1, p2):
        q1, q2 = _p(p1).reshape(-1), _p(p2).reshape(-1)
        ctx.save_for_backward(a, b, q1, q2)
        ctx.pshape = p1.shape
        return ops.affine2(a, b, q1, q2)

    @staticmethod
    def backward(ctx, g):
        a, b, q1, q2 = ctx.saved_tensors
        da, db, d1, d2 = ops.affine2_bwd(_cg(g), a, b, q1, q2)
        return da, db, d1.view(ctx.pshape), d2.view(ctx.pshape)


def affine2(a, b, p1, p2):
    a, b = _cg(a), _cg(b)
    if _needs_grad(a, b, p1, p2):
        return _Affine2.apply(a, b, p1, p2)
    return ops.affine2(a, b, _p(p1).reshape(-1), _p(p2).reshape(-1))


class _GatherRows(torch.autograd.Function):
    """y[j] = x[idx[j]]: SparseDispatcher.dispatch, ``inp[self._batch_index]`` for one expert (moce_ir.py:94-104)."""

    @staticmethod
    def forward(ctx, x, idx):
        ctx.save_for_backward(idx)
        ctx.shape = x.shape
        n = idx.numel()
        y = torch.empty((n,) + tuple(x.shape[1:]), device=x.device, dtype=x.dtype)
        return ops.rows_copy(x, y, n, src_idx=idx)

    @staticmethod
    def backward(ctx, g):
        (idx,) = ctx.saved_tensors
        g = _cg(g)
        dx = torch.zeros(ctx.shape, device=g.device, dtype=g.dtype)
        ops.rows_copy(g, dx, idx.numel(), dst_idx=idx)
        return dx, None


def gather_rows(x, idx):
    """x [B, C, H, W] (sample-contiguous), idx int32 [n] -> [n, C, H, W]"""
    x = x.contiguous()
    return _GatherRows.apply(x, idx) if _needs_grad(x) else ops.rows_copy(
        x, torch.empty((idx.numel(),) + tuple(x.shape[1:]), device=x.device, dtype=x.dtype), idx.numel(), src_idx=idx)


class _Combine(torch.autograd.Function):
    """out[b] = sum over experts e and their samples j with idx_e[j] == b of g_e[j] * y_e[j]: SparseDispatcher.combine,
    ``zeros.index_add(0, batch_index, stitched * nonzero_gates)`` (moce_ir.py:106-123), experts in ascending order."""

    @staticmethod
    def forward(ctx, B, idxs, *ts):
        ne = len(idxs)
        ys, gs = ts[:ne], ts[ne:]
        out = torch.zeros((B,) + tuple(ys[0].shape[1:]), device=ys[0].device, dtype=ys[0].dtype)
        for y, g, idx in zip(ys, gs, idxs):
            ops.rows_copy(y, out, idx.numel(), dst_idx=idx, scale=g, accumulate=True)
        ctx.save_for_backward(*ys, *gs, *idxs)
        ctx.ne = ne
        return out

    @staticmethod
    def backward(ctx, dout):
        ne = ctx.ne
        sv = ctx.saved_tensors
        ys, gs, idxs = sv[:ne], sv[ne:2 * ne], sv[2 * ne:]
        dout = _cg(dout).to(ys[0].dtype)
        dys, dgs = [], []
        for y, g, idx in zip(ys, gs, idxs):
            n = idx.numel()
            dys.append(ops.rows_copy(dout, torch.empty_like(y), n, src_idx=idx, scale=g))
            dgs.append(ops.rowdot(y, dout, n, b_idx=idx))
        return (None, None) + tuple(dys) + tuple(dgs)


def combine_rows(B, ys, gs, idxs):
    """ys[e] [n_e, C, H, W] (contiguous), gs[e] fp32 [n_e], idxs[e] int32 [n_e] -> [B, C, H, W]"""
    ys = [y.contiguous() for y in ys]
    gs = [g.float().contiguous() for g in gs]
    return _Combine.apply(B, tuple(idxs), *ys, *gs)
