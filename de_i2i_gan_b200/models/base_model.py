"""Mirror of models/base_model.py:7-87: network discovery by the ``net`` attribute prefix, init / save / load glue and
the loss dispatcher ``_cal_loss`` - whose four loss types run the fused reduction kernels of libdei2i.so (loss value
and d(loss)/d(input) in one pass) instead of torch.nn.functional."""
import torch

from .. import functional as Fn
from .networks import load_network, save_network


class BaseModel:
    def __init__(self, opt):
        self.opt = opt
        self.network_prefix = "net"

    @property
    def networks(self):
        """{'G': netG, 'D': netD, ...}: every nn.Module attribute whose name starts with 'net' (base_model.py:18-36)."""
        return {name.replace(self.network_prefix, ""): mod for name, mod in self.__dict__.items()
                if name.startswith(self.network_prefix) and isinstance(mod, torch.nn.Module)}

    def init_weights(self):
        """base_model.py:38-45; networks whose short name ends with '_' are skipped."""
        print(f"initialize model's parameters using {self.opt.init_type} with variance={self.opt.init_variance}")
        for name, net in self.networks.items():
            if not name.endswith("_"):
                net.init_weights(self.opt.init_type, self.opt.init_variance)

    def save(self, epoch):
        for name, net in self.networks.items():
            save_network(net, name, epoch, self.opt)

    def load(self, epoch):
        print(f"load model's weights from epoch {epoch}")
        for name, net in self.networks.items():
            load_network(net, name, epoch, self.opt)

    def load_network(self, network_name, epoch):
        print(f"load net_{network_name}'s weights from epoch {epoch}")
        load_network(self.networks[network_name], network_name, epoch, self.opt)

    def __repr__(self):
        out = ""
        for name, net in self.networks.items():
            bar = "=" * 50 + f"{self.network_prefix + name:^8}" + "=" * 50 + "\n"
            out += bar + repr(net) + "\n" + bar
        return out

    def _cal_loss(self, logits, targets, loss_type):
        """base_model.py:68-80 - inputs of 'bce'/'cce' are unnormalised logits; every reduction is 'mean'."""
        if loss_type in ("bce", "bce_logits"):
            return Fn.bce_with_logits(logits, targets)
        if loss_type in ("cce", "cce_logits"):
            return Fn.cross_entropy_soft(logits, targets)
        if loss_type == "l1":
            return Fn.l1_loss(logits, targets)
        if loss_type in ("l2", "mse"):
            return Fn.mse_loss(logits, targets)
        raise ValueError(f"loss_type: {loss_type} is invalid")

    def update_per_epoch(self, epoch):
        for net in self.networks.values():
            if hasattr(net, "update_per_epoch"):
                net.update_per_epoch(epoch)
