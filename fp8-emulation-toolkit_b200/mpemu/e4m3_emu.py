"""E4M3Emulator — PTQ policy (reference: mpemu/e4m3_emu.py:23, defaults :46-49, inference :181-185)."""
from .emulator_base import EmulatorBase
from .qutils import TensorQuantConfig


class E4M3Emulator(EmulatorBase):
    tag = "e4m3_emulator"
    weight_rule = "conv_linear"

    def __init__(self, model, optimizer, sparse_config=None, device="cuda", train=False, verbose=False,
                 tensor_stats=False):
        super().__init__(model, optimizer, sparse_config, device=device, verbose=verbose, tensor_stats=tensor_stats)
        self.is_training = train

    def set_training_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("e4m3", "rne")
        self.wt_qconfig = TensorQuantConfig("e4m3", "rne")
        self.iact_qconfig = TensorQuantConfig("e4m3", "rne")
        self.oact_qconfig = None

    def set_default_inference_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("e4m3", "rne", "per-channel")
        self.wt_qconfig = TensorQuantConfig("e4m3", "rne", "per-channel")
        self.iact_qconfig = TensorQuantConfig("e4m3", "rne", "per-tensor")
        self.oact_qconfig = None
