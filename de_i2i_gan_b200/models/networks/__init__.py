"""Checkpoint glue with the reference's file format (models/networks/__init__.py:4-23): plain
``torch.save(net.state_dict())`` files, loader strips 'spade_'/'sean_' substrings and ignores 'mlp_latent' keys."""
import torch


def save_network(net, net_label, epoch, opt):
    save_dir = opt.ckpt_dir / opt.name
    save_dir.mkdir(parents=True, exist_ok=True)
    torch.save(net.state_dict(), save_dir / f"{epoch}_net_{net_label}.pth")


def load_network(net, net_label, epoch, opt):
    path = opt.ckpt_dir / opt.load_model_name / f"{epoch}_net_{net_label}.pth"
    weights = torch.load(path, map_location="cpu")
    weights = {k.replace("spade_", "").replace("sean_", ""): v for k, v in weights.items() if "mlp_latent" not in k}
    net.load_state_dict(weights, strict=False)
    return net.to(opt.device, non_blocking=True)
