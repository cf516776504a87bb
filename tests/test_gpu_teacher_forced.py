"""Teacher-forced module parity at the BASELINE batch sizes: EVERY hooked module of ResNet-50 (configs[2], PTQ E4M3,
batch 64 x 3 x 224 x 224) and of BERT-base (configs[3], hybrid training step, batch 12 x 384) is driven in isolation —
its forward-pre hook (weights + inputs), forward hook (output) and, in training, the output-gradient / weight-gradient
tensor hooks and the full-backward hook — on synthetic tensors that the reference run (CPU) and this run (GPU)
generate bit-identically (tests/golden/teacher.py).  Goldens: tests/golden/teacher_{rn50,bert}.npz, produced by
tests/golden/make_golden_teacher.py from the reference's own qutils / mpt_emu over its unmodified CPU op.

Tolerance: NONE for deterministic casts — every digest (two 64-bit checksums over all elements + 62 sampled bit
patterns) must be equal, and the sequence of (module, hook, dtype, scheme, scaling, in-place, shape) cast requests
must be the same list.  Stochastic casts (E5M2 gradients: the generators differ by design) are checked for the
properties of the rounding instead: every output is a member of the E5M2 grid, less than one E5M2 step away from its
input, and the mean rounding error over the tensor is within 5 standard errors of zero."""
import json
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import models  # noqa: E402
import teacher  # noqa: E402


@pytest.fixture(scope="module")
def mp():
    import mpemu
    from mpemu import qutils
    assert "fp8-emulation-toolkit_b200" in mpemu.__file__
    qutils.set_unscale_mode("fp32")          # the goldens come from the reference CPU op (fp32 un-scale)
    yield mpemu
    qutils.set_unscale_mode("fp16")


def load(tag):
    g = np.load(os.path.join(HERE, "golden", "teacher_%s.npz" % tag))
    meta = json.loads(str(g["meta"][0]))
    raw = json.loads(str(g["shapes"][0]))
    shapes = {k: ([(tuple(s), d) for s, d in v[:-1]], tuple(v[-1]) if v[-1] is not None else None) for k, v in raw.items()}
    digests = {k[2:]: g[k] for k in g.files if k.startswith("d|")}
    return meta, [str(n) for n in g["names"]], shapes, digests, [str(s) for s in g["log"]]


def compare(digests, want, log, want_log):
    assert log == want_log, "cast requests differ: first difference %r" % next(
        ((a, b) for a, b in zip(log, want_log) if a != b), (len(log), len(want_log)))
    assert sorted(digests) == sorted(want)
    bad = [k for k in want if not np.array_equal(digests[k], want[k])]
    assert not bad, "%d of %d results differ from the reference, first: %s" % (len(bad), len(want), bad[:5])


def test_resnet50_ptq_every_hooked_module_bit_exact(mp):
    torchvision = pytest.importorskip("torchvision")
    from mpemu import mpt_emu, qutils
    meta, names, shapes, want, want_log = load("rn50")
    torch.manual_seed(0)
    rn = torchvision.models.resnet50(weights=None).eval()
    ck = models.checksum(list(rn.parameters()))
    if abs(ck - meta["weight_checksum"]) > 1e-6 * abs(meta["weight_checksum"]):
        pytest.skip("random init differs from the fixture's (different torch build)")
    rn = rn.cuda()
    rn, emu = mpt_emu.quantize_model(rn, dtype="e4m3", list_exempt_layers=["conv1", "bn1", "fc"], device="cuda")
    assert list(emu.emulator.model_qconfig_dict.keys()) == names
    assert meta["batch"] == 64
    digests, log = teacher.drive(rn, qutils, names, shapes, "cuda", training=False)
    assert len(want) == 3 * 52                       # weight, input, output of each of the 52 hooked convolutions
    compare(digests, want, log, want_log)


def e5m2_stochastic_properties(role, src, out, cfg):
    assert cfg.dtype == "e5m2", role
    import fpemu_b200 as F
    x, q = src.float().reshape(-1), out.float().reshape(-1)
    assert torch.equal(F.cast(q.contiguous(), "E5M2_RTZ"), q), role + ": output off the E5M2 grid"
    h = x.half().float()                             # the op rounds to fp16 first (RNE), then stochastically to E5M2
    step = torch.exp2(torch.floor(torch.log2(h.abs().clamp_min(2.0 ** -14))) - 2)      # E5M2 spacing at |h|
    assert bool(((q - h).abs() < step * 1.0001).all()), role + ": moved by a full step or more"
    assert bool((torch.sign(q) * torch.sign(h) >= 0).all()), role
    err = (q - h).double()
    assert abs(err.mean().item()) <= 5.0 * err.std().item() / np.sqrt(err.numel()) + 1e-12, role + ": biased"


def test_bert_base_hybrid_every_hooked_module(mp):
    transformers = pytest.importorskip("transformers")
    from transformers import BertConfig, BertForQuestionAnswering
    from mpemu import mpt_emu, qutils
    meta, names, shapes, want, want_log = load("bert")
    torch.manual_seed(0)
    cfg = BertConfig()
    cfg.hidden_dropout_prob = 0.0
    cfg.attention_probs_dropout_prob = 0.0
    model = BertForQuestionAnswering(cfg).train()
    ck = models.checksum(list(model.parameters()))
    if abs(ck - meta["weight_checksum"]) > 1e-6 * abs(meta["weight_checksum"]):
        pytest.skip("random init differs from the fixture's (different torch / transformers build)")
    model = model.cuda()
    model, emu = mpt_emu.initialize(model, None, training_algo="hybrid", list_exempt_layers=["qa_outputs"], device="cuda")
    assert list(emu.emulator.model_qconfig_dict.keys()) == names
    assert (meta["batch"], meta["seq"]) == (12, 384)
    seen = []

    def on_stochastic(role, src, out, qc):
        seen.append(role)
        if len(seen) % 7 == 1:                       # every 7th gradient cast gets the full property check
            e5m2_stochastic_properties(role, src, out, qc)
    qutils.set_batched_hooks(False)                  # per-tensor hook behaviour: every cast goes through quantize_tensor
    try:
        digests, log = teacher.drive(model, qutils, names, shapes, "cuda", training=True, on_stochastic=on_stochastic)
    finally:
        qutils.set_batched_hooks(True)
    compare(digests, want, log, want_log)
    assert len(seen) == sum("stochastic" in s for s in want_log) == 97 + 97 + 72    # ograd, igrad, wtgrad


def test_synthetic_tensors_are_device_independent():
    for shape, seed, ls in (((1000,), 5, 0), ((64, 3, 7, 7), 123456, 2), ((12, 384, 16), 99, -9)):
        a = teacher.synth(shape, seed, "cpu", ls)
        b = teacher.synth(shape, seed, "cuda", ls).cpu()
        assert torch.equal(a.view(torch.int32), b.view(torch.int32))
        assert np.array_equal(teacher.digest(a), teacher.digest(b.cuda()))
    assert torch.equal(teacher.synth_ids((50, 3), 4, "cpu", 30522), teacher.synth_ids((50, 3), 4, "cuda", 30522).cpu())
