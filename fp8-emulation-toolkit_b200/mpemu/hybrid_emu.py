"""HybridEmulator — forward tensors E4M3 RNE per-tensor, gradients E5M2 stochastic
(reference: mpemu/hybrid_emu.py:24, defaults :56-61, calibration / inference :400-413)."""
from .emulator_base import EmulatorBase
from .qutils import TensorQuantConfig


class HybridEmulator(EmulatorBase):
    tag = "hybrid_emulator"
    hooks_layernorm = True

    def __init__(self, model, optimizer, sparse_config=None, device="cuda", verbose=False, tensor_stats=False):
        super().__init__(model, optimizer, sparse_config, device=device, verbose=verbose, tensor_stats=tensor_stats)
        if self.is_training and optimizer is not None and not self.using_apex:
            # the reference refuses optimizers that did not pass through apex amp (hybrid_emu.py:69-79)
            raise RuntimeError("Failed to initialize, optimizer did not pass through amp.initialize. "
                               "Call amp.initilize(model, optimizer, opt_level='O2') before caling this interface.")

    def set_training_qconfig(self):
        self.emb_qconfig = None
        self.wt_qconfig = TensorQuantConfig("e4m3", "rne", "per-tensor")
        self.iact_qconfig = TensorQuantConfig("e4m3", "rne", "per-tensor")
        self.oact_qconfig = TensorQuantConfig("e4m3", "rne", "per-tensor")
        self.igrad_qconfig = TensorQuantConfig("e5m2", "stochastic")
        self.ograd_qconfig = TensorQuantConfig("e5m2", "stochastic")
        self.wtgrad_qconfig = TensorQuantConfig("e5m2", "stochastic")

    def set_default_inference_qconfig(self):
        self.emb_qconfig = TensorQuantConfig("e3m4", "rne", "per-channel")
        self.wt_qconfig = TensorQuantConfig("e3m4", "rne", "per-channel")
        self.iact_qconfig = TensorQuantConfig("e4m3", "rne", "per-tensor")
        self.oact_qconfig = self.igrad_qconfig = self.ograd_qconfig = self.wtgrad_qconfig = None
