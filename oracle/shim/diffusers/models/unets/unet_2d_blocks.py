import torch


class UNetMidBlock2D(torch.nn.Module):
    def __init__(self, *a, **k):
        super().__init__()
