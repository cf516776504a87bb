"""E3M4Emulator — PTQ policy (reference: mpemu/e3m4_emu.py:22, defaults :45-48, calibration :159-162,
inference :165-168)."""
from .emulator_base import EmulatorBase
from .qutils import TensorQuantConfig


class E3M4Emulator(EmulatorBase):
    tag = "e3m4_emulator"
    weight_rule = "conv_linear"

    def __init__(self, model, optimizer, sparse_config=None, device="cuda", train=False, verbose=False,
                 tensor_stats=False):
        super().__init__(model, optimizer, sparse_config, device=device, verbose=verbose, tensor_stats=tensor_stats)
        self.is_training = train

    def set_training_qconfig(self):
        self.set_default_inference_qconfig()

    def set_calibration_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("e3m4", "rne", "per-tensor")
        self.wt_qconfig = TensorQuantConfig("e3m4", "rne", "per-tensor")
        self.iact_qconfig = TensorQuantConfig("e3m4", "rne", "per-tensor")
        self.oact_qconfig = None

    def set_default_inference_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("e3m4", "rne", "per-channel")
        self.wt_qconfig = TensorQuantConfig("e3m4", "rne", "per-channel")
        self.iact_qconfig = TensorQuantConfig("e3m4", "rne", "per-tensor")
        self.oact_qconfig = None
