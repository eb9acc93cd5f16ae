import torch.nn as nn
from torch.nn import init


class BaseNetwork(nn.Module):
    """Mirror of models/networks/base_network.py:5-65 (device property, parameter report, weight init)."""

    @staticmethod
    def modify_commandline_options(parser, is_train):
        return parser

    @property
    def device(self):
        return next(self.parameters()).device

    def print_network(self):
        n = sum(p.numel() for p in self.parameters())
        print(f"Network [{type(self).__name__}] was created. Total number of parameters: {n / 1e6:.1f} million. "
              "To see the architecture, do print(network).")

    def init_weights(self, init_type="normal", gain=0.02):
        """base_network.py:27-61: BatchNorm2d weight ~ N(1, gain), bias 0; Conv / Linear weight by ``init_type``
        (for spectral-normed layers ``m.weight`` aliases ``weight_orig``'s storage, so that is what gets drawn),
        bias 0; visited in ``nn.Module.apply`` order so the RNG stream matches the reference for the same seed."""
        def fill(m):
            cls = m.__class__.__name__
            if "BatchNorm2d" in cls:
                if getattr(m, "weight", None) is not None:
                    init.normal_(m.weight.data, 1.0, gain)
                if getattr(m, "bias", None) is not None:
                    init.constant_(m.bias.data, 0.0)
            elif hasattr(m, "weight") and ("Conv" in cls or "Linear" in cls):
                w = m.weight.data
                if init_type == "normal":
                    init.normal_(w, 0.0, gain)
                elif init_type == "xavier":
                    init.xavier_normal_(w, gain=gain)
                elif init_type == "xavier_uniform":
                    init.xavier_uniform_(w, gain=1.0)
                elif init_type == "kaiming":
                    init.kaiming_normal_(w, a=0, mode="fan_in")
                elif init_type == "orthogonal":
                    init.orthogonal_(w, gain=gain)
                elif init_type == "none":
                    m.reset_parameters()
                else:
                    raise NotImplementedError(f"initialization method [{init_type}] is not implemented")
                if getattr(m, "bias", None) is not None:
                    init.constant_(m.bias.data, 0.0)

        self.apply(fill)
        for child in self.children():
            if hasattr(child, "init_weights"):
                child.init_weights(init_type, gain)

    def update_per_epoch(self, epoch):
        pass
