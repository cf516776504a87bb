"""qutils — tensor / module quantization configs, scale derivation and the module hooks.

Same public names and call signatures as the reference's mpemu/qutils.py (cited per item), re-built around the
device-side kernels:
  * per-tensor `max` scaling: amax kernel -> scale kernel -> cast kernel, no host sync, no temporaries
    (reference: abs() temp + max + `.item()`-style sync, qutils.py:282-287);
  * per-channel scaling: ONE row-wise launch (reference: a Python loop of per-channel op calls, :289-311);
  * fine-grained groups: one row-wise launch per tensor over (channel x group) rows (:313-337).
Out of scope here: INT quantization (`quantize_to_integer`), tensor bin-dumps, TensorBoard stats,
sparsity, op patching with the CPU hardware model.
"""
import copy
from collections import OrderedDict

import torch

import fpemu_b200 as _F
from fpemu_b200 import _lib as _L
from .pytquant import fpemu_cuda

# "fp16": un-scale through half like the reference CUDA op (default); "fp32": like the reference CPU op.
_UNSCALE_FP32 = False


def set_unscale_mode(mode):
    """'fp16' (reference CUDA op semantics, default) or 'fp32' (reference CPU op semantics)."""
    global _UNSCALE_FP32
    if mode not in ("fp16", "fp32"):
        raise ValueError("unscale mode must be 'fp16' or 'fp32'")
    _UNSCALE_FP32 = (mode == "fp32")


class TensorQuantConfig(object):
    """Reference: TensorQuantConfig, mpemu/qutils.py:22-132 (same constructor, same attributes)."""
    dtypes_all = ["e5m2", "e4m3", "e3m4", "fp4", "bfloat16", "float16"]
    e5m2_modes = ["rtz", "stochastic", "rne", "rnaz", "rntz", "rpinf", "rninf",
                  "daz_stochastic", "daz_rne", "daz_rnaz", "daz_rntz"]
    e4m3_modes = ["rne", "stochastic"]
    e4m3_ieee_modes = ["ieee_rne", "ieee_stochastic"]
    e4m3_all_modes = e4m3_modes + e4m3_ieee_modes + ["rtz"]
    e3m4_modes = ["rne", "stochastic", "rtz"]
    fp4_modes = ["nearest"]
    bfloat16_modes = ["rne", "stochastic", "rtz"]
    float16_modes = ["rne", "stochastic", "daz_rne"]

    def __init__(self, dtype=None, scheme=None, scaling="None", group_size=1):
        self.dtype = dtype
        self.scheme = scheme
        parts = scaling.split("-")
        self.scaling = "tensor" in parts
        self.per_channel = "channel" in parts
        self.scaling_method = "mean" if "mean" in parts else "max"
        if "int" in self.dtype:
            self.scaling_method = "none"
        self.fine_grained = (scaling == "fine-grained")
        self.group_size = group_size if self.fine_grained else 1
        self.is_enabled = True
        self.check_validity(dtype, scheme)

    def check_validity(self, dtype, scheme):
        if "int" in dtype:
            return
        assert dtype in self.dtypes_all, "Invalid data type {}, supported: {}".format(dtype, self.dtypes_all)
        table = {"e5m2": self.e5m2_modes, "e4m3": self.e4m3_all_modes, "e3m4": self.e3m4_modes,
                 "fp4": self.fp4_modes, "bfloat16": self.bfloat16_modes, "float16": self.float16_modes}
        assert scheme in table[dtype], "scheme {} is not in {}".format(scheme, table[dtype])

    def get_flt_max(self):
        """Reference: qutils.py:100-114."""
        if self.dtype == "e5m2":
            return 57344.0
        if self.dtype == "e4m3":
            return 240.0 if self.scheme in self.e4m3_ieee_modes else 448.0
        if self.dtype == "e3m4":
            return 30.0
        if self.dtype == "fp4":
            return 1.0
        return None

    def get_flt_min(self):
        """Reference: qutils.py:116-128."""
        return {"e5m2": 1.5258789e-05, "e4m3": 1.9531250e-03, "e3m4": 1.5625000e-02,
                "fp4": 0.000244140625}.get(self.dtype)

    def mode_string(self):
        mode = self.dtype.upper()
        if self.scheme is not None:
            mode += "_" + self.scheme.upper()
        return mode

    def __repr__(self):
        how = ("per-channel" if self.per_channel else "per-tensor" if self.scaling
               else "fine-grained" if self.fine_grained else "None")
        return "[{}, scale: {}, method: {}]".format(self.dtype + "_" + str(self.scheme), how, self.scaling_method)


class ModuleQuantConfig(object):
    """Reference: ModuleQuantConfig, mpemu/qutils.py:137-198."""

    def __init__(self, wt_qconfig=None, iact_qconfig=None, oact_qconfig=None,
                 wtgrad_qconfig=None, igrad_qconfig=None, ograd_qconfig=None, patch_ops=False):
        self.wt_qconfig = wt_qconfig
        self.iact_qconfig = iact_qconfig
        self.oact_qconfig = oact_qconfig
        self.wtgrad_qconfig = wtgrad_qconfig
        self.igrad_qconfig = igrad_qconfig
        self.ograd_qconfig = ograd_qconfig
        self.device = 'cuda'
        self.patch_ops = patch_ops
        self.patch_impl = "None"
        self.sparse_config = None
        self.global_step = -1
        self.bin_dump = False
        self.tensor_stats = False
        self.tb_writer = None
        self.weight_master = None

    def register_output_grad_hook(self, output):
        """Output-gradient tensor hook (reference :159-171): out-of-place cast of the incoming gradient."""
        def quantize_output_grad(grad):
            if self.ograd_qconfig is not None and self.ograd_qconfig.is_enabled:
                grad = quantize_tensor(grad, self.ograd_qconfig, None, inplace=False)
            return grad
        if output.requires_grad:
            output.register_hook(quantize_output_grad)

    def register_weight_grad_hook(self, module):
        """Weight-gradient tensor hook (reference :173-189): in-place cast."""
        def quantize_weight_grad(grad):
            qc = self.wtgrad_qconfig
            if qc is not None and qc.is_enabled:
                if _GRADS.defer(module, qc, grad):
                    return grad              # cast in place, with every other deferred gradient, when this backward ends
                grad.data = quantize_tensor(grad.data, qc, None, inplace=True)
            return grad
        if type(module) in (torch.nn.Conv2d, torch.nn.Linear):
            if module.weight is not None and module.weight.requires_grad:
                module.weight.register_hook(quantize_weight_grad)

    def __repr__(self):
        if self.igrad_qconfig is None:
            return "[weights: {}, inputs: {}, output: {}]".format(self.wt_qconfig, self.iact_qconfig, self.oact_qconfig)
        return "[weights: {}, inputs: {}, output: {}, weight_grad: {}, input_grad: {}, output_grad: {}]".format(
            self.wt_qconfig, self.iact_qconfig, self.oact_qconfig, self.wtgrad_qconfig, self.igrad_qconfig,
            self.ograd_qconfig)


class ModuleQuantParams(object):
    """Reference: mpemu/qutils.py:204-214 (stateless formats carry no parameters)."""

    def __init__(self, wt_qparams=None, iact_qparams=None, oact_qparams=None,
                 wtgrad_qparams=None, igrad_qparams=None, ograd_qparams=None):
        self.wt_qparams = wt_qparams
        self.iact_qparams = iact_qparams
        self.oact_qparams = oact_qparams
        self.wtgrad_qparams = wtgrad_qparams
        self.igrad_qparams = igrad_qparams
        self.ograd_qparams = ograd_qparams


def quantize_to_integer(tensor, mode, inplace=False):
    raise NotImplementedError("INT quantization (reference qutils.py:216-247) is outside the fpemu cast path")


def fpemu_device_fn(tensor, mode, inplace=True, scale=1.0):
    """Reference: mpemu/qutils.py:249-261.  CUDA tensors only — there is no CPU op here."""
    if "INT" in mode:
        return quantize_to_integer(tensor, mode.split("_")[0], inplace=inplace)
    if not tensor.is_cuda:
        raise RuntimeError("fpemu: the B200 build has no CPU path; move the tensor to a CUDA device")
    if not (tensor.requires_grad and torch.is_grad_enabled()) and not tensor.is_sparse and mode != 'NONE' \
            and tensor.dtype != torch.float64:
        # nothing to record for autograd (weights' .data, gradients, inference): skip the Function wrapper
        src = tensor if tensor.is_contiguous() else tensor.contiguous()
        out = src if (inplace and src is tensor) else torch.empty_like(src)
        if isinstance(scale, torch.Tensor) and scale.is_cuda and scale.dtype == torch.float32 and scale.numel() == 1:
            _F.cast(src, mode, out=out, d_scale=scale, unscale_fp32=_UNSCALE_FP32)
        else:
            _F.cast(src, mode, out=out, scale=float(scale), unscale_fp32=_UNSCALE_FP32)
        if inplace and src is not tensor:
            tensor.copy_(out)
            return tensor
        return out
    return fpemu_cuda.FPEmuOp.apply(tensor, mode, inplace, scale)


def _mean_scale(mean, flt_min, floor):
    """`mean` scaling (reference :276-281 / :294-301): scale = max(1, flt_min / mean) with a floor on the mean.
    The reference divides a Python float by a float32 tensor, i.e. in float32 — so does this."""
    import numpy as np
    mean = abs(float(mean))
    if not mean > floor:
        return 1.0                       # `mean` becomes the Python float flt_min there: flt_min / flt_min
    # `flt_min / abs(mean)` with `mean` a 0-dim tensor is `mean.reciprocal() * flt_min` (torch/_tensor.py __rdiv__)
    scale = float(np.float32(np.float32(1.0) / np.float32(mean)) * np.float32(flt_min))
    return 1.0 if scale < 1.0 else scale


def _scaled_cast_nograd(tensor, mode, inplace, kind, flt_max, rows, cols, clamp):
    src = tensor if tensor.is_contiguous() else tensor.contiguous()
    out = src if (inplace and src is tensor) else torch.empty_like(src)
    if kind == "tensor":
        _F.cast_amax_scaled(src, mode, out=out, flt_max=flt_max, unscale_fp32=_UNSCALE_FP32)
    else:
        _F.cast_rowwise(src, mode, out=out, flt_max=flt_max, clamp=clamp, rows=rows, cols=cols,
                        unscale_fp32=_UNSCALE_FP32)
    if inplace and src is not tensor:
        tensor.copy_(out)
        return tensor
    return out


def _scaled_cast(tensor, *args):
    """Autograd wrapper only where a graph is being recorded (activations in training)."""
    if tensor.requires_grad and torch.is_grad_enabled():
        return _ScaledCast.apply(tensor, *args)
    return _scaled_cast_nograd(tensor, *args)


class _ScaledCast(torch.autograd.Function):
    """Device-side scaled casts with the same straight-through gradient as FPEmuOp."""

    @staticmethod
    def forward(ctx, tensor, mode, inplace, kind, flt_max, rows, cols, clamp):
        src = tensor if tensor.is_contiguous() else tensor.contiguous()
        out = src if (inplace and src is tensor) else torch.empty_like(src)
        if kind == "tensor":
            _F.cast_amax_scaled(src, mode, out=out, flt_max=flt_max, unscale_fp32=_UNSCALE_FP32)
        else:
            _F.cast_rowwise(src, mode, out=out, flt_max=flt_max, clamp=clamp, rows=rows, cols=cols,
                            unscale_fp32=_UNSCALE_FP32)
        if inplace:
            if src is not tensor:
                tensor.copy_(out)
            ctx.mark_dirty(tensor)
            return tensor
        return out

    @staticmethod
    def backward(ctx, grad):
        return grad, None, None, None, None, None, None, None


def _rows_mode_scaled(tensor, mode, inplace, rows, cols, flt_min):
    """`mean` scaling over dim-0 channels or fine-grained groups: the reference takes the *mode* of each row
    (qutils.py:297, :321) and calls the op once per row.  Same numbers from ONE torch.mode over dim 1 (checked equal to
    the per-row calls, ties included), the scale expression of :298-301 on the device, and one row-wise launch with
    those scales: no .item()."""
    src = tensor.detach()
    src = src if src.is_contiguous() else src.contiguous()
    m = torch.abs(torch.mode(src.reshape(rows, cols), dim=1).values).float()
    m = torch.where(m > 1e-6, m, torch.full_like(m, flt_min))
    scales = torch.clamp(flt_min / m, min=1.0).contiguous()
    out = src if (inplace and src.data_ptr() == tensor.data_ptr()) else torch.empty_like(src)
    _F.cast_rows_scaled(src, mode, scales, out=out, rows=rows, cols=cols, unscale_fp32=_UNSCALE_FP32)
    if inplace and out.data_ptr() != tensor.data_ptr():
        tensor.copy_(out)
        return tensor
    return tensor if inplace else out


def _supported_dtype(t):
    return t.dtype in (torch.float32, torch.float16, torch.bfloat16)


def _quantize_double(tensor, qtconfig, qtparams, mode, inplace):
    """float64 tensors (rare: the reference dispatches them, cuda/fpemu_kernels.cu:1059).  Unscaled and per-tensor scaled
    casts run the double kernel (fpemu_cast_f64) with the scale derived exactly as the reference derives it — torch
    arithmetic in double on the device (`flt_max / vmax` = reciprocal * flt_max, qutils.py:282-285; the mean rule
    :276-281), narrowed to float where the op's `float scale` argument narrows it — and no host sync.  Per-channel and
    fine-grained modes cast a float copy (their scales are derived in float there)."""
    fmt = _L.parse_mode(mode)[0]
    takes_scale = fmt not in (_L.FMT_BF16, _L.FMT_FP16)
    if qtconfig.scaling is True or not (qtconfig.per_channel or qtconfig.fine_grained):
        scale = 1.0
        if qtconfig.scaling is True and takes_scale and tensor.numel() > 0:
            flat = tensor.detach().flatten()
            if qtconfig.scaling_method == "mean":
                flt_min = qtconfig.get_flt_min()
                mean = torch.abs(torch.mean(torch.abs(flat)))
                mean = torch.where(mean > 1e-5, mean, torch.full_like(mean, flt_min))
                scale = torch.clamp(flt_min / mean, min=1.0)
            else:
                scale = qtconfig.get_flt_max() / torch.max(torch.abs(flat))
                scale = torch.where(scale > 3.275e+04, torch.full_like(scale, 6.55e+04), scale)
            scale = scale.float().reshape(1)
        return fpemu_device_fn(tensor, mode, inplace=inplace, scale=scale)
    q = quantize_tensor(tensor.float(), qtconfig, qtparams, inplace=False).to(torch.float64)
    if inplace:
        tensor.copy_(q)
        return tensor
    return q


def quantize_tensor(tensor, qtconfig, qtparams, inplace=False):
    """Single entry point for quantizing a tensor — reference: mpemu/qutils.py:264-345, same branches:
    per-tensor (max / mean), per-channel over dim 0, fine-grained groups, unscaled."""
    qtconfig.check_validity(qtconfig.dtype, qtconfig.scheme)
    mode = qtconfig.mode_string()
    if "INT" in mode:
        return quantize_to_integer(tensor, mode.split("_")[0], inplace=inplace)
    if not tensor.is_cuda:
        raise RuntimeError("fpemu: the B200 build has no CPU path; move the tensor to a CUDA device")
    if tensor.dtype == torch.float64:
        return _quantize_double(tensor, qtconfig, qtparams, mode, inplace)
    fmt = _L.parse_mode(mode)[0]
    takes_scale = fmt not in (_L.FMT_BF16, _L.FMT_FP16)
    flt_max = qtconfig.get_flt_max()

    if qtconfig.scaling is True:
        if qtconfig.scaling_method == "mean":
            # same expression as the reference (:276-281), kept on the device: no .item(), no host sync
            flt_min = qtconfig.get_flt_min()
            if takes_scale and _supported_dtype(tensor) and tensor.is_contiguous() and tensor.numel() > 0:
                scale = _F.mean_scale(tensor.detach(), flt_min, 1e-5)        # two small kernels, deterministic
            else:
                mean = torch.abs(torch.mean(torch.abs(tensor.detach().flatten())))
                mean = torch.where(mean > 1e-5, mean, torch.full_like(mean, flt_min))
                scale = torch.clamp(flt_min / mean, min=1.0).float().reshape(1)
            return fpemu_device_fn(tensor, mode, inplace=inplace, scale=scale)
        if takes_scale and _supported_dtype(tensor):
            return _scaled_cast(tensor, mode, inplace, "tensor", flt_max, 0, 0, True)
        return fpemu_device_fn(tensor, mode, inplace=inplace)

    if qtconfig.per_channel is True:
        channels = tensor.shape[0]
        cols = tensor.numel() // max(channels, 1)
        if qtconfig.scaling_method == "mean":
            # the reference takes the *mode* of each channel here (:297) and calls the op once per channel.  Same
            # numbers from ONE torch.mode over dim 1 (checked equal to the per-channel calls, ties included), the
            # scale expression of :298-301 on the device, and one row-wise launch with those scales: no .item().
            flt_min = qtconfig.get_flt_min()
            if takes_scale and _supported_dtype(tensor) and channels > 0 and cols > 0:     # incl. E4M3_IEEE, FP4
                return _rows_mode_scaled(tensor, mode, inplace, channels, cols, flt_min)
            out = tensor if inplace else torch.zeros_like(tensor)
            for c in range(channels):
                sub = tensor.select(0, c).detach()
                m = abs(torch.mode(torch.flatten(sub)).values.item())
                scale = _mean_scale(m, flt_min, 1e-6)
                out.select(0, c).data.copy_(fpemu_device_fn(sub, mode, inplace=False, scale=scale))
            return out
        if takes_scale and _supported_dtype(tensor):
            return _scaled_cast(tensor, mode, inplace, "rows", flt_max, channels, cols, True)
        return fpemu_device_fn(tensor, mode, inplace=inplace)

    if qtconfig.fine_grained is True:
        # groups of `group_size` input channels inside every output channel (:313-337): with C divisible into
        # equal chunks these are contiguous runs -> rows of a (K * chunks) x run matrix; no scale clamp here
        k = tensor.shape[0]
        in_ch = tensor.shape[1] if tensor.dim() > 1 else 1
        chunks = max(1, int(in_ch / qtconfig.group_size))
        if (qtconfig.scaling_method == "max" and takes_scale and _supported_dtype(tensor)
                and tensor.dim() > 1 and in_ch % chunks == 0):
            rows = k * chunks
            return _scaled_cast(tensor, mode, inplace, "rows", flt_max, rows, tensor.numel() // rows, False)
        if (qtconfig.scaling_method == "mean" and takes_scale and _supported_dtype(tensor)
                and tensor.dim() > 1 and in_ch % chunks == 0 and tensor.numel() > 0):
            rows = k * chunks                               # the same groups, `mean` (torch.mode) scaling: one launch
            return _rows_mode_scaled(tensor, mode, inplace, rows, tensor.numel() // rows, qtconfig.get_flt_min())
        out = tensor if inplace else tensor.clone()
        for kk in range(k):
            crs = out.select(0, kk)
            for sub in crs.chunk(chunks, 0):
                if qtconfig.scaling_method == "mean":
                    m = abs(torch.mode(torch.flatten(sub)).values.item())
                    scale = _mean_scale(m, qtconfig.get_flt_min(), 1e-6)
                else:
                    # a 0-dim tensor like the reference's (`reciprocal * flt_max`); an all-zero group gives inf
                    scale = (flt_max / torch.max(torch.abs(sub))).float().reshape(1)
                sub.copy_(fpemu_device_fn(sub.contiguous(), mode, inplace=False, scale=scale))
        return out

    return fpemu_device_fn(tensor, mode, inplace=inplace)


# --------------------------------------------------------------------------------------------- model plumbing
def get_or_update_model_quant_config_dict(model, filter_module_types, mod_qconfig, prefix="",
                                          model_qconfig_dict=None, override=False, exempt_modules=[]):
    """Reference: mpemu/qutils.py:351-365."""
    if model_qconfig_dict is None:
        model_qconfig_dict = OrderedDict()
    for name, module in model.named_modules():
        fname = prefix + name
        if type(module) in filter_module_types and fname not in exempt_modules:
            if fname not in model_qconfig_dict or override:
                model_qconfig_dict[fname] = copy.deepcopy(mod_qconfig)
        elif type(module) not in filter_module_types and fname in model_qconfig_dict:
            model_qconfig_dict.pop(fname)
    return model_qconfig_dict


def get_quantization_parameters(tensor, qtconfig):
    return None       # stateless formats (reference :445-453)


def quantize_weights_of_a_module(module, wt_qconfig, name=""):
    """Reference: mpemu/qutils.py:458-468 (cast into a new tensor, then copy back): same values from ONE in-place pass."""
    w = module.weight.data
    if w.is_cuda and w.is_contiguous():
        quantize_tensor(w, wt_qconfig, None, inplace=True)
    else:
        w.copy_(quantize_tensor(w, wt_qconfig, None))


def quantize_model_weights(model, model_qconfig_dict):
    """Reference: mpemu/qutils.py:470-475."""
    for name, module in model.named_modules():
        if name in model_qconfig_dict and model_qconfig_dict[name].wt_qconfig is not None:
            quantize_weights_of_a_module(module, model_qconfig_dict[name].wt_qconfig, name=name)


# ------------------------------------------------------------------------------------------ batched launches
# The reference casts tensor by tensor from Python.  Two groups of those casts do not depend on anything computed
# during the step and are issued here as ONE multi-tensor sweep each (fpemu_amax_multi + fpemu_cast_multi_amax /
# fpemu_cast_multi) instead of one launch per module:
#   * the weights of all hooked modules of a model in training (re-quantized in place every forward,
#     qutils.py:520-522): cast when the ROOT module's forward starts; each module's own pre-forward hook then finds
#     its weight done.  Same values as the per-module casts (weights do not change during a forward).
#   * the weight gradients (tensor hooks, qutils.py:173-189): a hook whose parameter has no accumulated gradient yet
#     leaves the incoming gradient untouched; when the backward pass ends, the `.grad` tensors of all such parameters
#     are cast in place in one launch.  With a gradient already accumulated (micro-batching) the hook casts at once,
#     as the reference does, because cast(a) + cast(b) != cast(a + b).
# Calling a hooked module on its own (no root forward) or set_batched_hooks(False) gives the per-tensor behaviour.
_BATCH_HOOKS = True


def set_batched_hooks(flag):
    global _BATCH_HOOKS
    _BATCH_HOOKS = bool(flag)


def _plan_for(cache, tensors):
    key = tuple((t.data_ptr(), t.numel()) for t in tensors)
    plan = cache.get(key)
    if plan is None:
        if len(cache) > 16:
            cache.clear()
        try:
            plan = _F.MultiTensorPlan(tensors)
        except ValueError:
            plan = False                                    # unaligned storage: per-tensor casts
        cache[key] = plan
    return plan


def _batchable(qc, t):
    return (qc is not None and qc.is_enabled and not qc.per_channel and not qc.fine_grained and t is not None
            and t.is_cuda and t.is_contiguous() and _supported_dtype(t) and t.numel() > 0
            and (qc.scaling is not True or qc.scaling_method == "max") and "int" not in qc.dtype)


def _cast_list(plans, tensors, qc):
    """In-place cast of a list of same-dtype tensors that share one TensorQuantConfig: one sweep, or False."""
    plan = _plan_for(plans, tensors)
    if plan is False:
        return False
    mode = qc.mode_string()
    if qc.scaling is True and _L.parse_mode(mode)[0] not in (_L.FMT_BF16, _L.FMT_FP16):
        plan.amax_scaled_cast(mode, flt_max=qc.get_flt_max(), unscale_fp32=_UNSCALE_FP32)
    else:
        plan.cast(mode, unscale_fp32=_UNSCALE_FP32)
    return True


class _WeightBatch(object):
    """Forward-pre hook of the ROOT module of a model in training: one sweep over the weights of all hooked modules."""

    def __init__(self, modules):
        self.modules = modules
        self.plans = {}
        self.active = False            # True only while the root module's forward is running
        self.gen = 0

    def done(self, root, inputs, output):
        self.active = False

    def __call__(self, root, inputs):
        self.active = False
        if not _BATCH_HOOKS or torch.cuda.is_current_stream_capturing():
            return None
        self.active = True
        self.gen += 1
        groups = {}
        for m in self.modules:
            qc = m.qconfig.wt_qconfig
            w = getattr(m, "weight", None)
            if w is None or not _batchable(qc, w.data) or qc.dtype in ["bfloat16"]:
                continue
            groups.setdefault((qc.mode_string(), qc.scaling, w.dtype, w.device), (qc, []))[1].append(m)
        for qc, mods in groups.values():
            if len(mods) > 1 and _cast_list(self.plans, [m.weight.data for m in mods], qc):
                for m in mods:
                    m._fpemu_weight_done = (self, self.gen)
        return None


class _GradBatch(object):
    """Weight gradients whose cast was deferred to the end of the running backward pass."""

    def __init__(self):
        self.pending = []
        self.plans = {}

    def defer(self, module, qc, grad):
        if not _BATCH_HOOKS or module.weight.grad is not None or not _batchable(qc, grad) \
                or torch.cuda.is_current_stream_capturing():
            return False
        # one callback per deferred gradient: the first one to run casts everything, the others find nothing to do
        # (a single callback per pass would be lost for good if a backward pass raised before running it)
        torch.autograd.Variable._execution_engine.queue_callback(self.flush)
        self.pending.append((module, qc))
        return True

    def flush(self):
        pending, self.pending = self.pending, []
        groups, seen = {}, set()
        for m, qc in pending:
            g = m.weight.grad
            if g is None or id(m) in seen:
                continue
            seen.add(id(m))
            if not g.is_contiguous():
                g.data = quantize_tensor(g.data, qc, None, inplace=True)
                continue
            groups.setdefault((qc.mode_string(), g.dtype, g.device), (qc, []))[1].append(g.data)
        for qc, grads in groups.values():
            if not _cast_list(self.plans, grads, qc):
                for g in grads:
                    quantize_tensor(g, qc, None, inplace=True)


_GRADS = _GradBatch()


def quantize_module_weights_and_inputs(self, input):
    """Forward-pre hook (reference :478-549): weights in place, inputs out of place."""
    qc = self.qconfig
    if qc.wt_qconfig is not None and qc.wt_qconfig.dtype in ["bfloat16"]:
        qc.weight_master = self.weight.data.clone()
    done = getattr(self, "_fpemu_weight_done", None)
    if done is not None and done[0].active and done[0].gen == done[1]:
        self._fpemu_weight_done = None                       # cast by the root module's sweep at the start of this forward
    elif qc.wt_qconfig is not None and qc.wt_qconfig.is_enabled:
        self.weight.data = quantize_tensor(self.weight.data, qc.wt_qconfig, self.qparams.wt_qparams, inplace=True)
    if qc.iact_qconfig is not None and qc.iact_qconfig.is_enabled:
        input = tuple(quantize_tensor(t, qc.iact_qconfig, None, inplace=False) if torch.is_tensor(t) and
                      t.is_floating_point() else t for t in input)
    return input


def quantize_module_gradient(self, grad_input, grad_output):
    """Full-backward hook (reference :551-584): input gradients out of place."""
    qc = self.qconfig
    if qc.igrad_qconfig is not None and qc.igrad_qconfig.is_enabled:
        grad_input = tuple(quantize_tensor(g, qc.igrad_qconfig, self.qparams.igrad_qparams, inplace=False)
                           if g is not None else None for g in grad_input)
    if qc.wt_qconfig is not None and qc.wt_qconfig.dtype in ["bfloat16"] and qc.weight_master is not None:
        self.weight.data = qc.weight_master.clone()
    return grad_input


_OUTPUT_CAST_OUT_OF_PLACE = False


def set_output_cast_out_of_place(flag):
    """False (default): module outputs are cast in place like the reference (`qutils.py:611`), bit-for-bit the
    reference's behaviour on this PyTorch, including its loss of the output-gradient hook on view outputs.
    True: while autograd is recording, outputs are cast into a new tensor (see quantize_module_output)."""
    global _OUTPUT_CAST_OUT_OF_PLACE
    _OUTPUT_CAST_OUT_OF_PLACE = bool(flag)


def quantize_module_output(self, input, output):
    """Forward hook (reference :586-616): register the output-gradient hook, cast the output in place."""
    qc = self.qconfig
    if torch.is_tensor(output):
        qc.register_output_grad_hook(output)
        if qc.oact_qconfig is not None and qc.oact_qconfig.is_enabled:
            # The reference always casts in place.  While autograd is recording, the output of a Linear on a 3-D
            # input is a VIEW of the 2-D addmm result; an in-place op on such a view makes autograd rebuild the graph
            # through CopySlices (four extra full-tensor copies per module in backward, 288 per BERT-base step) and
            # the output-gradient hook registered two lines above never fires, because the gradient no longer
            # passes through the node it was registered on.  set_output_cast_out_of_place(True) returns a new
            # tensor instead: same forward values, no copies, and the output-gradient hook fires.
            inplace = not (_OUTPUT_CAST_OUT_OF_PLACE and output.requires_grad and torch.is_grad_enabled())
            output = quantize_tensor(output, qc.oact_qconfig, self.qparams.oact_qparams, inplace=inplace)
    return output


def add_quantization_hooks(model, model_qconfig_dict, is_training=False):
    """Reference: mpemu/qutils.py:618-633."""
    handles = []
    for name, module in model.named_modules():
        if name in model_qconfig_dict:
            handles.append(module.register_forward_pre_hook(quantize_module_weights_and_inputs))
            handles.append(module.register_forward_hook(quantize_module_output))
            if is_training:
                if module.qconfig.igrad_qconfig is not None:
                    handles.append(module.register_full_backward_hook(quantize_module_gradient))
                module.qconfig.register_weight_grad_hook(module)
    if is_training:
        hooked = [m for name, m in model.named_modules() if name in model_qconfig_dict and hasattr(m, "weight")]
        if len(hooked) > 1:
            batch = _WeightBatch(hooked)
            handles.append(model.register_forward_pre_hook(batch, prepend=True))
            handles.append(model.register_forward_hook(batch.done, always_call=True))
    return handles


def reset_quantization_setup(model, model_qconfig_dict):
    """Reference: mpemu/qutils.py:636-642."""
    for name, module in model.named_modules():
        if name in model_qconfig_dict:
            module.name = name
            module.qconfig = copy.deepcopy(model_qconfig_dict[name])
            module.qparams = ModuleQuantParams()


def _set_flag(model, model_qconfig_dict, attr, value):
    for name, module in model.named_modules():
        if name in model_qconfig_dict:
            cfg = getattr(module.qconfig, attr)
            if cfg is not None:
                cfg.is_enabled = value


def set_quantize_weights_flag(model, model_qconfig_dict, set_value):
    _set_flag(model, model_qconfig_dict, "wt_qconfig", set_value)        # reference :644-648


def set_quantize_outputs_flag(model, model_qconfig_dict, set_value):
    _set_flag(model, model_qconfig_dict, "oact_qconfig", set_value)      # reference :650-654


def set_quantize_inputs_flag(model, model_qconfig_dict, set_value):
    _set_flag(model, model_qconfig_dict, "iact_qconfig", set_value)      # reference :656-660
