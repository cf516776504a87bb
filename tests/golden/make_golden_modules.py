#!/usr/bin/env python
"""Module-level golden outputs from the REFERENCE Python stack (imported from /root/reference, CPU).

    python oracle/build_ref.py cpu && python tests/golden/make_golden_modules.py

The reference package imports `fpemu_cpp` (its compiled CPU op) and the CPU hardware-model extensions at
package import; here `fpemu_cpp` is the unmodified build from oracle/_ref/ and the two hardware-model modules
(not on the cast path) are empty stand-ins.  Everything else is the reference's own code:
qutils.quantize_tensor, mpt_emu.quantize_model, mpt_emu.initialize.  Output: tests/golden/modules.npz.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from oracle import ref_loader  # noqa: E402
import models  # noqa: E402

# the build a real install of the reference produces on this host: setup.py compiles with -march=native, i.e. the
# AVX-512 paths (the scalar bf16 path mis-rounds, cpp/fpemu_impl.cpp:321-324; RNE / stochastic modes used below are
# the same in both builds for E5M2 / E4M3 / E3M4)
sys.modules["fpemu_cpp"] = ref_loader.load("fpemu_cpp_ref_avx512") or ref_loader.load("fpemu_cpp_ref_scalar")
for stub in ("simple_gemm_dev", "simple_conv2d_dev"):
    sys.modules[stub] = types.ModuleType(stub)
sys.path.insert(0, os.environ.get("FPEMU_REFERENCE_ROOT", "/root/reference"))
import mpemu as ref                                   # noqa: E402  (the reference package)
from mpemu import mpt_emu as ref_mpt                  # noqa: E402
from mpemu.qutils import TensorQuantConfig, quantize_tensor   # noqa: E402

assert ref.__file__.startswith("/root/reference"), ref.__file__
torch.set_num_threads(1)
out = {}

# ---- A. quantize_tensor branches (qutils.py:264-345)
g = torch.Generator().manual_seed(5)
w = torch.randn(24, 16, 3, 3, generator=g) * torch.exp2(torch.randint(-4, 4, (24, 1, 1, 1), generator=g).float())
a = torch.randn(4, 16, 8, 8, generator=g) * 3.0
out["qt_weight"] = w.numpy()
out["qt_act"] = a.numpy()
for dt, sch, scaling, gs, src in [("e4m3", "rne", "per-tensor", 1, "qt_act"), ("e4m3", "rne", "per-channel", 1, "qt_weight"),
                                  ("e3m4", "rne", "per-channel", 1, "qt_weight"), ("e5m2", "rne", "per-tensor", 1, "qt_act"),
                                  ("e4m3", "rne", "fine-grained", 4, "qt_weight"), ("e5m2", "rne", "None", 1, "qt_act"),
                                  ("e5m2", "rne", "per-tensor-mean", 1, "qt_act"), ("bfloat16", "rne", "None", 1, "qt_act")]:
    cfg = TensorQuantConfig(dt, sch, scaling, gs)
    x = torch.from_numpy(out[src]).clone()
    q = quantize_tensor(x, cfg, None, inplace=False)
    out["qt|%s|%s|%s|%d|%s" % (dt, sch, scaling, gs, src)] = q.detach().numpy().copy()

# ---- B. PTQ of a small CNN (mpt_emu.quantize_model, E4M3 per-channel weights + per-tensor inputs)
for dtype in ("e4m3", "e3m4"):
    m = models.make("cnn").eval()
    x = torch.randn(2, 3, 16, 16, generator=torch.Generator().manual_seed(9))
    out["cnn_input"] = x.numpy()
    out["cnn_fp32_out"] = m(x).detach().numpy()
    m, emu = ref_mpt.quantize_model(m, dtype=dtype, list_exempt_layers=["conv1"], device="cpu")
    with torch.no_grad():
        y = m(x)
    out["cnn_%s_out" % dtype] = y.numpy()
    for k, v in m.state_dict().items():
        if k.endswith("weight") and v.dim() > 1:
            out["cnn_%s_w|%s" % (dtype, k)] = v.numpy().copy()

# ---- B2. batch-norm folding + PTQ (fuse_bn=True path)
# (quantize_model(fuse_bn=True) crashes upstream — `mpt.emulator` is still None at mpt_emu.py:204 — so this is
#  the flow of examples/inference/classifier/imagenet_test.py:197-209: quantize, then fuse and re-prepare)
m = models.make("cnn").eval()
m, emu = ref_mpt.quantize_model(m, dtype="e4m3", list_exempt_layers=["conv1"], device="cpu")
m = emu.fuse_bnlayers_and_quantize_model(m)
with torch.no_grad():
    out["cnn_e4m3_fused_out"] = m(torch.from_numpy(out["cnn_input"])).numpy()

# ---- C. ResNet-50 PTQ, BASELINE config 3 at batch 2 (weights regenerated from the seed on the other side)
try:
    import torchvision
    torch.manual_seed(0)
    rn = torchvision.models.resnet50(weights=None).eval()
    out["rn50_weight_checksum"] = np.array([models.checksum(list(rn.parameters()))])
    xr = torch.randn(2, 3, 224, 224, generator=torch.Generator().manual_seed(1))
    out["rn50_input_checksum"] = np.array([models.checksum([xr])])
    with torch.no_grad():
        out["rn50_fp32_logits"] = rn(xr).numpy()
    rn, emu = ref_mpt.quantize_model(rn, dtype="e4m3", list_exempt_layers=["conv1", "bn1", "fc"], device="cpu")
    with torch.no_grad():
        out["rn50_e4m3_logits"] = rn(xr).numpy()
        # noise floor of this randomly initialised network: the reference's own logits under a 1e-7 input perturbation
        yp = rn(xr + 1e-7 * torch.randn(xr.shape, generator=torch.Generator().manual_seed(2))).numpy()
    out["rn50_sensitivity"] = np.array([np.linalg.norm(yp - out["rn50_e4m3_logits"]) / np.linalg.norm(out["rn50_e4m3_logits"])])
    out["rn50_layer1_0_conv1_w"] = rn.layer1[0].conv1.weight.detach().numpy().copy()
    out["rn50_hooked_modules"] = np.array([len(emu.emulator.model_qconfig_dict)])
except ImportError:
    print("torchvision missing: ResNet-50 fixture skipped")

# ---- D. hybrid training step (mpt_emu.initialize, optimizer=None: no apex here)
m = models.make("mlp").train()
ids = torch.randint(0, 100, (8, 12), generator=torch.Generator().manual_seed(3))
tgt = torch.randint(0, 4, (8,), generator=torch.Generator().manual_seed(4))
out["mlp_ids"] = ids.numpy()
out["mlp_tgt"] = tgt.numpy()
m, emu = ref_mpt.initialize(m, None, training_algo="hybrid", list_exempt_layers=["head"], device="cpu")
acts = {}
for nm, mod in m.named_modules():
    if nm:
        mod.register_forward_hook(lambda mod_, inp, outp, nm=nm: acts.__setitem__(nm, outp.detach().clone()))
logits = m(ids)
for nm, v in acts.items():
    out["mlp_act|" + nm] = v.numpy()
loss = torch.nn.functional.cross_entropy(logits, tgt)
loss.backward()
out["mlp_logits"] = logits.detach().numpy()
out["mlp_loss"] = np.array([loss.item()])
for k, p in m.named_parameters():
    out["mlp_grad|" + k] = p.grad.detach().numpy().copy()
    out["mlp_w_after|" + k] = p.detach().numpy().copy()    # weights are quantized in place during training

np.savez_compressed(os.path.join(HERE, "modules.npz"), **out)
print("modules.npz written:", len(out), "arrays")
