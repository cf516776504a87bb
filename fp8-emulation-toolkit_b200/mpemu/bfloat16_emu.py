"""Bfloat16Emulator — everything BFLOAT16 RNE (reference: mpemu/bfloat16_emu.py:27, defaults :55-61,
inference :105-111).  The reference's master-parameter code calls a module that does not exist
(`quantemu_cuda.QuantEmuOp`, :125-176); that dead path is not reproduced."""
from .emulator_base import EmulatorBase
from .qutils import TensorQuantConfig


class Bfloat16Emulator(EmulatorBase):
    tag = "bfloat16_emulator"

    def set_training_qconfig(self):
        self.emb_qconfig = None
        for role in ("wt", "iact", "oact", "igrad", "ograd", "wtgrad"):
            setattr(self, role + "_qconfig", TensorQuantConfig("bfloat16", "rne"))

    def set_default_inference_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("bfloat16", "rne")
        self.wt_qconfig = TensorQuantConfig("bfloat16", "rne")
        self.iact_qconfig = TensorQuantConfig("bfloat16", "rne")
        self.oact_qconfig = self.igrad_qconfig = self.ograd_qconfig = self.wtgrad_qconfig = None

    def update_master_params(self, optimizer):
        return
