import torch


class _Inert(torch.nn.Module):
    def __init__(self, *a, **k):
        super().__init__()


class ResnetBlock2D(_Inert):
    pass


class Downsample2D(_Inert):
    pass


class Upsample2D(_Inert):
    pass
