"""Dense ResNet head (reference ``language_table/train/networks/dense_resnet.py:24-73``)."""
import torch.nn as nn
import torch.nn.functional as F


def dense(cin: int, cout: int) -> nn.Linear:
    """``nn.Dense`` with the reference's normal(0.05) kernel and bias initialisers."""
    lin = nn.Linear(cin, cout)
    nn.init.normal_(lin.weight, std=0.05)
    nn.init.normal_(lin.bias, std=0.05)
    return lin


class ResnetDenseBlock(nn.Module):
    def __init__(self, width: int):
        super().__init__()
        self.fc1, self.fc2, self.fc3 = dense(width, width // 4), dense(width // 4, width // 4), dense(width // 4, width)

    def forward(self, x):
        y = self.fc1(F.relu(x))
        y = self.fc2(F.relu(y))
        y = self.fc3(F.relu(y))
        return x + y


class DenseResnet(nn.Module):
    def __init__(self, in_features: int, width: int, num_blocks: int, value_net: bool = False):
        super().__init__()
        self.inp = dense(in_features, width)
        self.blocks = nn.ModuleList([ResnetDenseBlock(width) for _ in range(num_blocks)])
        self.value = dense(width, 1) if value_net else None

    def forward(self, x):
        x = self.inp(x)
        for b in self.blocks:
            x = b(x)
        return self.value(x) if self.value is not None else x
