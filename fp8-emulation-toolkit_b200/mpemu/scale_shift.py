"""ScaleShift: an eval-mode BatchNorm2d folded into one multiply-add per channel
(reference: mpemu/scale_shift.py:12-73; needed by the ResNet-50 PTQ flow, not a quantization scale)."""
import torch


class ScaleShift(torch.nn.Module):
    def __init__(self, num_features):
        super().__init__()
        self.num_features = num_features
        self.weight = torch.nn.Parameter(torch.ones(num_features))
        self.bias = torch.nn.Parameter(torch.zeros(num_features))

    def forward(self, x):
        # channel dimension is 1: move it last, broadcast, move it back
        return (x.transpose(1, -1) * self.weight + self.bias).transpose(1, -1)

    def __repr__(self):
        return "ScaleShift({})".format(self.num_features)

    @staticmethod
    def generate_from_batchnorm(bn):
        """scale = gamma / sqrt(var + eps), shift = beta - gamma * mean / sqrt(var + eps)."""
        ss = ScaleShift(bn.num_features)
        with torch.no_grad():
            inv_std = torch.sqrt(bn.running_var + bn.eps)
            gamma = bn.weight if bn.weight is not None else torch.ones_like(bn.running_var)
            beta = bn.bias if bn.bias is not None else torch.zeros_like(bn.running_var)
            ss.weight.copy_(gamma.div(inv_std))
            ss.bias.copy_(beta - gamma.mul(bn.running_mean).div(inv_std))
        return ss


def replace_batchnorms_with_scaleshifts(model):
    if isinstance(model, torch.nn.BatchNorm2d):
        return ScaleShift.generate_from_batchnorm(model)
    for name, child in model.named_children():
        setattr(model, name, replace_batchnorms_with_scaleshifts(child))
    return model
