#!/usr/bin/env python
"""Teacher-forced per-module goldens from the REFERENCE Python stack (imported from /root/reference, CPU).

    python oracle/build_ref.py cpu && python tests/golden/make_golden_teacher.py [rn50] [bert]

For BASELINE configs[2] (ResNet-50 PTQ, E4M3, batch 64 x 3 x 224 x 224) and configs[3] (BERT-base hybrid training step,
batch 12 x 384) every hooked module of the reference-prepared model is driven in isolation by tests/golden/teacher.py:
its hook functions (mpemu/qutils.py:478-616 and the tensor hooks :159-189) are called on integer-hash synthetic tensors
of the shapes the module sees at that batch size, and every deterministic result is stored as a digest.  The reference
package runs over its own unmodified CPU op (scalar build, oracle/_ref/), exactly as in make_golden_modules.py.
Output: tests/golden/teacher_rn50.npz, tests/golden/teacher_bert.npz (a few hundred KB)."""
import copy
import json
import os
import sys
import time
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from oracle import ref_loader  # noqa: E402
import models  # noqa: E402
import teacher  # noqa: E402

SCALAR = ref_loader.load("fpemu_cpp_ref_scalar")
assert SCALAR is not None, "run oracle/build_ref.py cpu first"
sys.modules["fpemu_cpp"] = SCALAR
for stub in ("simple_gemm_dev", "simple_conv2d_dev"):
    sys.modules[stub] = types.ModuleType(stub)
sys.path.insert(0, os.environ.get("FPEMU_REFERENCE_ROOT", "/root/reference"))
import mpemu as ref                                   # noqa: E402  (the reference package)
from mpemu import mpt_emu as ref_mpt                  # noqa: E402
from mpemu import qutils as ref_qutils                # noqa: E402

assert ref.__file__.startswith("/root/reference"), ref.__file__


def save(tag, meta, names, shapes, digests, log):
    out = {"meta": np.array([json.dumps(meta)]), "names": np.array(names),
           "shapes": np.array([json.dumps({k: [[list(s), d] for s, d in v[0]] + [list(v[1]) if v[1] else None]
                                           for k, v in shapes.items()})]),
           "log": np.array(log)}
    for k, v in digests.items():
        out["d|" + k] = v
    np.savez_compressed(os.path.join(HERE, "teacher_%s.npz" % tag), **out)
    print("teacher_%s.npz: %d modules, %d digests, %d casts logged" % (tag, len(names), len(digests), len(log)))


def rn50(batch=64):
    import torchvision
    torch.manual_seed(0)
    rn = torchvision.models.resnet50(weights=None).eval()
    ck = models.checksum(list(rn.parameters()))
    plain = copy.deepcopy(rn)
    rn, emu = ref_mpt.quantize_model(rn, dtype="e4m3", list_exempt_layers=["conv1", "bn1", "fc"], device="cpu")
    names = list(emu.emulator.model_qconfig_dict.keys())
    t0 = time.time()
    shapes = teacher.record_shapes(plain, names, lambda m: m(torch.zeros(batch, 3, 224, 224)))
    del plain
    digests, log = teacher.drive(rn, ref_qutils, names, shapes, "cpu", training=False)
    print("rn50 driven in %.0f s" % (time.time() - t0))
    save("rn50", {"model": "torchvision resnet50, seed 0", "batch": batch, "weight_checksum": ck,
                  "call": "quantize_model(dtype='e4m3', list_exempt_layers=['conv1','bn1','fc'])"}, names, shapes, digests, log)


def bert(batch=12, seq=384):
    from transformers import BertConfig, BertForQuestionAnswering
    torch.manual_seed(0)
    cfg = BertConfig()
    cfg.hidden_dropout_prob = 0.0
    cfg.attention_probs_dropout_prob = 0.0
    model = BertForQuestionAnswering(cfg).train()
    ck = models.checksum(list(model.parameters()))
    plain = copy.deepcopy(model)
    model, emu = ref_mpt.initialize(model, None, training_algo="hybrid", list_exempt_layers=["qa_outputs"], device="cpu")
    names = list(emu.emulator.model_qconfig_dict.keys())
    ids = teacher.synth_ids((batch, seq), 1, "cpu", cfg.vocab_size)
    t0 = time.time()
    shapes = teacher.record_shapes(plain, names, lambda m: m(input_ids=ids))
    del plain
    digests, log = teacher.drive(model, ref_qutils, names, shapes, "cpu", training=True)
    print("bert driven in %.0f s" % (time.time() - t0))
    save("bert", {"model": "transformers BertForQuestionAnswering(BertConfig()), dropout 0, seed 0", "batch": batch, "seq": seq,
                  "weight_checksum": ck,
                  "call": "initialize(model, None, training_algo='hybrid', list_exempt_layers=['qa_outputs'])"},
         names, shapes, digests, log)


if __name__ == "__main__":
    torch.set_num_threads(max(1, (os.cpu_count() or 2) - 1))
    want = sys.argv[1:] or ["rn50", "bert"]
    if "rn50" in want:
        rn50()
    if "bert" in want:
        bert()
