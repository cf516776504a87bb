"""E5M2Emulator — the "direct" policy (reference: mpemu/e5m2_emu.py:24, defaults :56-61, inference :407-413)."""
from .emulator_base import EmulatorBase
from .qutils import TensorQuantConfig


class E5M2Emulator(EmulatorBase):
    tag = "e5m2_emulator"
    hooks_layernorm = True

    def set_training_qconfig(self):
        self.emb_qconfig = None
        self.wt_qconfig = TensorQuantConfig("e5m2", "rne")
        self.iact_qconfig = TensorQuantConfig("e5m2", "rne")
        self.oact_qconfig = TensorQuantConfig("e5m2", "rne")
        self.igrad_qconfig = TensorQuantConfig("e5m2", "stochastic")
        self.ograd_qconfig = TensorQuantConfig("e5m2", "stochastic")
        self.wtgrad_qconfig = TensorQuantConfig("e5m2", "stochastic")

    def set_calibration_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("e5m2", "rne")
        self.wt_qconfig = TensorQuantConfig("e5m2", "rne")
        self.iact_qconfig = TensorQuantConfig("e5m2", "rne")
        self.oact_qconfig = None

    def set_default_inference_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("e5m2", "rne", "per-channel-mean")
        self.wt_qconfig = TensorQuantConfig("e5m2", "rne", "per-channel-mean")
        self.iact_qconfig = TensorQuantConfig("e5m2", "rne")
        self.oact_qconfig = self.igrad_qconfig = self.ograd_qconfig = self.wtgrad_qconfig = None
