"""Small deterministic models shared by the golden generator (reference stack, CPU) and the GPU tests (this
repo's mpemu).  Weights come from torch.manual_seed, so both sides build identical parameters."""
import torch


class TinyCNN(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = torch.nn.Conv2d(3, 16, 3, padding=1)
        self.bn1 = torch.nn.BatchNorm2d(16)
        self.conv2 = torch.nn.Conv2d(16, 32, 3, padding=1)
        self.conv3 = torch.nn.Conv2d(32, 32, 1)
        self.pool = torch.nn.AdaptiveAvgPool2d(1)
        self.fc = torch.nn.Linear(32, 10)

    def forward(self, x):
        x = torch.relu(self.bn1(self.conv1(x)))
        x = torch.relu(self.conv2(x))
        x = torch.relu(self.conv3(x))
        return self.fc(torch.flatten(self.pool(x), 1))


class TinyMLP(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.embeddings = torch.nn.Embedding(100, 32)
        self.fc1 = torch.nn.Linear(32, 64)
        self.norm = torch.nn.LayerNorm(64)
        self.fc2 = torch.nn.Linear(64, 32)
        self.head = torch.nn.Linear(32, 4)

    def forward(self, ids):
        x = self.embeddings(ids)
        x = torch.nn.functional.gelu(self.norm(self.fc1(x)))
        x = self.fc2(x).mean(dim=1)
        return self.head(x)


def make(name, seed=0):
    torch.manual_seed(seed)
    m = {"cnn": TinyCNN, "mlp": TinyMLP}[name]()
    if name == "cnn":     # non-trivial BN statistics
        with torch.no_grad():
            m.bn1.running_mean.copy_(torch.randn(16) * 0.1)
            m.bn1.running_var.copy_(torch.rand(16) + 0.5)
    return m


def checksum(tensors):
    """Order-sensitive float64 checksum used to make sure both sides really built the same tensors."""
    s = 0.0
    for i, t in enumerate(tensors):
        s += float(t.detach().double().sum().item()) * (1.0 + 1e-3 * (i % 7)) + float(t.detach().double().abs().sum().item())
    return s
