"""mpt_emu — the two user entry points, same signatures as the reference:
`initialize(model, optimizer, training_algo, ...)` (mpemu/mpt_emu.py:146-191) and
`quantize_model(model, ..., dtype, ...)` (:193-231), plus the `MPTEmulator` façade (:29-120)."""
from .bfloat16_emu import Bfloat16Emulator
from .e3m4_emu import E3M4Emulator
from .e4m3_emu import E4M3Emulator
from .e5m2_emu import E5M2Emulator
from .hybrid_emu import HybridEmulator


class MPTEmulator(object):
    valid_training_methods = ["direct", "hybrid"]
    valid_hw_patches = ["none", "simple"]
    valid_pruning_methods = ["none", "fine-grained", "unstructured", "adaptive", "auto"]

    def __init__(self, device='cuda', training_algo='direct', hw_patch='none', pruning_algo='none'):
        if training_algo.lower() not in self.valid_training_methods:
            raise RuntimeError("mpt_emulator: the requested training method {} is not supported, use one of the "
                               "following: {}".format(training_algo, self.valid_training_methods))
        if hw_patch.lower() not in self.valid_hw_patches:
            raise RuntimeError("mpt_emulator: the requested hardware emulation {} is not supported, use one of the "
                               "following: {}".format(hw_patch, self.valid_hw_patches))
        if pruning_algo.lower() not in self.valid_pruning_methods:
            raise RuntimeError("mpt_emulator: the requested pruning method {} is not supported, use one of the "
                               "following: {}".format(pruning_algo, self.valid_pruning_methods))
        if pruning_algo.lower() != 'none':
            raise NotImplementedError("mpt_emulator: pruning is outside the fpemu cast path (DESIGN.md §7)")
        self.device = device
        self.training_algo = training_algo.lower()
        self.hw_patch = hw_patch.lower()
        self.pruning_method = 'none'
        self.emulator = None
        self.sparse_config = None

    def blacklist_modules(self, list_modules):
        self.emulator.blacklist_modules(list_modules)

    def whitelist_modules(self, list_modules):
        self.emulator.whitelist_modules(list_modules)

    def optimizer_step(self, optimizer):
        self.emulator.optimizer_step(optimizer)

    def update_global_steps(self, global_steps):
        self.emulator.global_steps = global_steps

    def set_master_param_precision(self, master_params):
        self.emulator.set_master_param_precision(master_params)

    def set_embedding_precision(self, emb_precision, emb_norm=False):
        self.emulator.set_embedding_precision(emb_precision, emb_norm)

    def fuse_bnlayers_and_quantize_model(self, model):
        self.emulator.is_training = False
        return self.emulator.fuse_layers_and_quantize_model(model)

    def set_calibration_qconfig(self):
        self.emulator.set_calibration_qconfig()

    def set_default_inference_qconfig(self):
        self.emulator.set_default_inference_qconfig()

    def __repr__(self):
        return repr(self.emulator)


def _check_device(device, hw_patch):
    if device != 'cuda' and not str(device).startswith('cuda'):
        raise RuntimeError("mpt_emulator: this build runs the cast on CUDA devices only (no CPU path)")
    if hw_patch.lower() != 'none':
        raise RuntimeError("mpt_emulator: HW patching ops is only alowed on 'cpu' device.")


def initialize(model, optimizer, training_algo='direct', hw_patch='none', pruning_algo='none',
               list_exempt_layers=None, list_layers_output_fused=None, device="cuda", verbose=False):
    if model is None:
        raise RuntimeError("mpt_emulator: Undefined model and optimizer, call this after model and optimizer are "
                           "initilized.")
    _check_device(device, hw_patch)
    mpt = MPTEmulator(device=device, training_algo=training_algo, hw_patch=hw_patch, pruning_algo=pruning_algo)
    cls = HybridEmulator if mpt.training_algo == 'hybrid' else E5M2Emulator
    mpt.emulator = cls(model, optimizer, None, device=device, verbose=verbose)
    mpt.emulator.set_master_param_precision("fp16")
    mpt.emulator.enable_hw_patching(mpt.hw_patch.upper())
    mpt.emulator.prepare_model(model, list_exempt_layers, list_layers_output_fused)
    if verbose:
        mpt.emulator.print_config()
    return model, mpt


def quantize_model(model, optimizer=None, dtype="none", calibrate=False, hw_patch="none", fuse_bn=False,
                   list_exempt_layers=None, list_layers_output_fused=None, device="cuda", verbose=False):
    if model is None:
        raise RuntimeError("mpt_emulator: Undefined model , call this after model is initilized.")
    _check_device(device, hw_patch)
    table = {'E5M2': E5M2Emulator, 'E4M3': E4M3Emulator, 'E3M4': E3M4Emulator, 'HYBRID': HybridEmulator,
             'BFLOAT16': Bfloat16Emulator}
    if dtype.upper() not in table:
        raise RuntimeError("mpt_emulator: the requested data type {} is not supported, use one of the following: {}"
                           .format(dtype, sorted(table)))
    mpt = MPTEmulator(device=device, hw_patch=hw_patch)
    was_training = model.training
    model.train(False)                      # emulators read model.training; PTQ is an inference flow
    mpt.emulator = table[dtype.upper()](model, optimizer, None, device=device, verbose=verbose)
    model.train(was_training)
    mpt.emulator.is_training = False
    if fuse_bn:
        mpt.emulator.list_exempt_layers = list_exempt_layers
        mpt.emulator.list_layers_output_fused = list_layers_output_fused
        model = mpt.emulator.fuse_layers_and_quantize_model(model)
        return model, mpt
    if calibrate:
        mpt.set_calibration_qconfig()
    else:
        mpt.set_default_inference_qconfig()
    mpt.emulator.enable_hw_patching(mpt.hw_patch.upper())
    mpt.emulator.prepare_model(model, list_exempt_layers, list_layers_output_fused)
    if verbose:
        mpt.emulator.print_config()
    return model, mpt
