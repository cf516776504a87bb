"""Teacher-forced, module-by-module comparison of the hook layer.

End-to-end logits of a randomly initialised deep network amplify a single rounding flip, so they cannot show whether
the hook on layer 40 casts the right tensor with the right configuration.  Here every hooked module is driven in
ISOLATION with inputs both sides can reproduce exactly: integer-hash pseudo-random tensors (int64 arithmetic and
int -> float conversions are identical on CPU and CUDA) of the shapes the module sees at the BASELINE batch sizes.
The module's own hook functions are then called directly — forward-pre (weights + inputs), forward (output), the
output-gradient / weight-gradient tensor hooks and the full-backward hook — and each result is reduced to a digest
(two 64-bit checksums of the bit patterns + 62 sampled elements).  `quantize_tensor` is wrapped to log which
configuration every cast was asked for.

The SAME driver runs in two places:
  * tests/golden/make_golden_teacher.py: over the REFERENCE package (its qutils, its CPU op)  -> tests/golden/teacher_*.npz
  * tests/test_gpu_teacher_forced.py:    over this repo's mpemu on the GPU                    -> compared bit for bit
Deterministic casts must agree bit for bit; stochastic casts (different generators by design) must have been asked
for with the same configuration, and are property-checked on the GPU side."""
import zlib

import numpy as np
import torch

NSAMPLES = 62


def _hash32(idx, seed):
    h = (idx * 2654435761 + (seed * 40503 + 12345)) & 0xFFFFFFFF
    h = h ^ (h >> 15)
    h = (h * 2246822519) & 0xFFFFFFFF
    h = h ^ (h >> 13)
    h = (h * 3266489917) & 0xFFFFFFFF
    return h ^ (h >> 16)


def synth(shape, seed, device, log2_scale=0):
    """float32 tensor of `shape`: (k / 65536 - 0.5) * 2^(e - 4 + log2_scale), k in [0, 65536), e in [0, 8) from an
    integer hash of the element index — every step is exact, so CPU and CUDA produce the same bits."""
    n = int(np.prod(shape)) if len(shape) else 1
    h = _hash32(torch.arange(n, dtype=torch.int64, device=device), int(seed))
    mant = (h & 0xFFFF).to(torch.float32) / 65536.0 - 0.5
    e = ((h >> 16) & 7).to(torch.int32) - 4 + int(log2_scale)
    pow2 = ((e + 127) << 23).view(torch.float32)             # 2^e built from its encoding: no pow() involved
    return (mant * pow2).reshape(shape)


def synth_ids(shape, seed, device, modulo):
    n = int(np.prod(shape)) if len(shape) else 1
    h = _hash32(torch.arange(n, dtype=torch.int64, device=device), int(seed))
    return (h % int(modulo)).reshape(shape)


def name_seed(name, k=0):
    return (zlib.crc32(name.encode()) + 7919 * k) & 0x7FFFFFFF


def digest(t):
    """int64[2 + NSAMPLES]: sum of the bit patterns, index-weighted sum (both modulo 2^64), and NSAMPLES sampled bit
    patterns.  NaN payloads are canonicalised (they differ between x86 and CUDA conversions)."""
    x = t.detach().contiguous().reshape(-1).to(torch.float32)
    bits = x.view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    bits = torch.where(torch.isnan(x), torch.full_like(bits, 0x7FC00000), bits)
    n = bits.numel()
    w = (torch.arange(n, dtype=torch.int64, device=bits.device) % 65521) + 1
    pick = (torch.arange(NSAMPLES, dtype=torch.int64, device=bits.device) * max(n - 1, 0)) // (NSAMPLES - 1)
    out = torch.cat([bits.sum().reshape(1), (bits * w).sum().reshape(1), bits[pick] if n else bits.new_zeros(NSAMPLES)])
    return out.cpu().numpy().astype(np.int64)


class Spy(object):
    """Wraps `qutils.quantize_tensor`, logging (role, dtype, scheme, scaling, inplace, shape) per call."""

    def __init__(self, qutils, on_stochastic=None):
        self.qutils = qutils
        self.orig = qutils.quantize_tensor
        self.role = ""
        self.log = []
        self.on_stochastic = on_stochastic   # callable(role, input clone, output, qtconfig): property checks (GPU side)

    def __enter__(self):
        def wrapped(tensor, qtconfig, qtparams, inplace=False):
            how = ("per-channel" if qtconfig.per_channel else "per-tensor" if qtconfig.scaling is True
                   else "fine-grained" if qtconfig.fine_grained else "none")
            self.log.append("%s|%s|%s|%s|%s|%s|%s" % (self.role, qtconfig.dtype, qtconfig.scheme, how,
                                                        qtconfig.scaling_method, int(bool(inplace)),
                                                        "x".join(str(int(s)) for s in tensor.shape)))
            keep = self.on_stochastic is not None and "stochastic" in str(qtconfig.scheme)
            src = tensor.detach().clone() if keep else None
            out = self.orig(tensor, qtconfig, qtparams, inplace=inplace)
            if keep:
                self.on_stochastic(self.role, src, out.detach(), qtconfig)
            return out
        self.qutils.quantize_tensor = wrapped
        return self

    def __exit__(self, *exc):
        self.qutils.quantize_tensor = self.orig


def record_shapes(model, names, example_fn):
    """Input / output shapes of the named modules during one plain forward of `model` (no emulation hooks yet)."""
    shapes, handles = {}, []
    mods = dict(model.named_modules())

    def make(nm):
        def hook(mod, inp, out):
            if nm in shapes:
                return
            ins = [(tuple(t.shape), str(t.dtype)) for t in inp if torch.is_tensor(t)]
            shapes[nm] = (ins, tuple(out.shape) if torch.is_tensor(out) else None)
        return hook
    for nm in names:
        handles.append(mods[nm].register_forward_hook(make(nm)))
    with torch.no_grad():
        example_fn(model)
    for h in handles:
        h.remove()
    return shapes


def drive(model, qutils, hooked_names, shapes, device, training, on_stochastic=None):
    """Calls the hook functions of every hooked module on synthetic tensors.  Returns (digests: {key: int64 array},
    log: [str]) — digests only for deterministic results."""
    mods = dict(model.named_modules())
    digests = {}
    with Spy(qutils, on_stochastic) as spy:
        for nm in hooked_names:
            if nm not in shapes:
                continue
            mod = mods[nm]
            ins, out_shape = shapes[nm]
            inputs = []
            for i, (shp, dt) in enumerate(ins):
                if "int" in dt:
                    modulo = getattr(mod, "num_embeddings", 1000)
                    inputs.append(synth_ids(shp, name_seed(nm, i), device, modulo))
                else:
                    inputs.append(synth(shp, name_seed(nm, i), device, log2_scale=1))
            spy.role = nm + ":pre"
            new_in = qutils.quantize_module_weights_and_inputs(mod, tuple(inputs))
            if getattr(mod, "weight", None) is not None:
                digests[nm + "|weight"] = digest(mod.weight.data)
            for i, t in enumerate(new_in):
                if torch.is_tensor(t) and t.is_floating_point():
                    digests[nm + "|in%d" % i] = digest(t)
            if out_shape is None:
                continue
            out = synth(out_shape, name_seed(nm, 50), device, log2_scale=2)
            if training:
                out = out.requires_grad_() * 1.0          # a non-leaf that requires grad, like a real module output
            spy.role = nm + ":out"
            out_q = qutils.quantize_module_output(mod, new_in, out)
            oq = mod.qconfig.oact_qconfig
            if oq is None or "stochastic" not in str(oq.scheme):
                digests[nm + "|out"] = digest(out_q)
            if not training:
                continue
            g_out = synth(out_shape, name_seed(nm, 60), device, log2_scale=-9)
            spy.role = nm + ":ograd"
            if out_q.requires_grad:
                out_q.backward(g_out)                     # fires the output-gradient tensor hook, if it survived
            g_ins = tuple(synth(shp, name_seed(nm, 70 + i), device, log2_scale=-9) for i, (shp, dt) in enumerate(ins)
                          if "int" not in dt)
            spy.role = nm + ":igrad"
            if mod.qconfig.igrad_qconfig is not None and len(g_ins):
                res = qutils.quantize_module_gradient(mod, g_ins, (g_out,))
                ig = mod.qconfig.igrad_qconfig
                if "stochastic" not in str(ig.scheme):
                    for i, t in enumerate(res):
                        digests[nm + "|igrad%d" % i] = digest(t)
            w = getattr(mod, "weight", None)
            if w is not None and w.requires_grad:
                spy.role = nm + ":wtgrad"
                w.grad = None
                w.backward(synth(tuple(w.shape), name_seed(nm, 80), device, log2_scale=-9))
                wg = mod.qconfig.wtgrad_qconfig
                if wg is not None and "stochastic" not in str(wg.scheme) and w.grad is not None:
                    digests[nm + "|wtgrad"] = digest(w.grad)
                w.grad = None
    return digests, list(spy.log)
