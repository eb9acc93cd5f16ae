"""TEST INFRASTRUCTURE ONLY - plain-torch restatement of DiffAugment (defectGAN/utils/diffaug.py:9-75) with EXPLICIT random
parameters, so that the CUDA kernels, this oracle and the reference can be fed the same draws.

``draw(x, policy)`` makes the reference's RNG calls (same shapes, dtypes, order: diffaug.py:22-58); ``apply(x, policy, drawn)``
evaluates the ops with a different formulation than the reference's gather / scatter code (slicing for the translation, index
comparisons for the cutout window).  Pinned against the unmodified reference in tests/test_oracle_vs_reference.py.
"""
import torch


def draw(x, policy):
    n, _, h, w = x.shape
    out = {}
    for name in policy.split(","):
        if name == "color":  # diffaug.py:22-36: three torch.rand(N,1,1,1)
            out[name] = dict(brightness=torch.rand(n, 1, 1, 1, dtype=x.dtype, device=x.device),
                             saturation=torch.rand(n, 1, 1, 1, dtype=x.dtype, device=x.device),
                             contrast=torch.rand(n, 1, 1, 1, dtype=x.dtype, device=x.device))
        elif name == "translation":  # diffaug.py:39-42
            sx, sy = int(h * 0.125 + 0.5), int(w * 0.125 + 0.5)
            out[name] = dict(tx=torch.randint(-sx, sx + 1, size=[n, 1, 1], device=x.device),
                             ty=torch.randint(-sy, sy + 1, size=[n, 1, 1], device=x.device))
        elif name == "cutout":  # diffaug.py:55-58
            ch, cw = int(h * 0.5 + 0.5), int(w * 0.5 + 0.5)
            out[name] = dict(ox=torch.randint(0, h + (1 - ch % 2), size=[n, 1, 1], device=x.device),
                             oy=torch.randint(0, w + (1 - cw % 2), size=[n, 1, 1], device=x.device))
        else:
            raise KeyError(name)
    return out


def apply(x, policy, drawn):
    n, c, h, w = x.shape
    for name in policy.split(","):
        d = drawn[name]
        if name == "color":
            x = x + (d["brightness"] - 0.5)                                   # diffaug.py:22-24
            m = x.mean(dim=1, keepdim=True)
            x = (x - m) * (d["saturation"] * 2) + m                           # diffaug.py:27-30
            m = x.mean(dim=(1, 2, 3), keepdim=True)
            x = (x - m) * (d["contrast"] + 0.5) + m                           # diffaug.py:33-36
        elif name == "translation":                                           # diffaug.py:39-52: y[i, j] = x[i + tx, j + ty], zeros outside
            out = []
            for k in range(n):
                tx, ty = int(d["tx"][k]), int(d["ty"][k])
                y = torch.zeros_like(x[k])
                i0, i1 = max(0, -tx), min(h, h - tx)
                j0, j1 = max(0, -ty), min(w, w - ty)
                if i1 > i0 and j1 > j0:
                    y[:, i0:i1, j0:j1] = x[k][:, i0 + tx:i1 + tx, j0 + ty:j1 + ty]
                out.append(y)
            x = torch.stack(out)
        elif name == "cutout":                                                # diffaug.py:55-71
            ch, cw = int(h * 0.5 + 0.5), int(w * 0.5 + 0.5)
            ii = torch.arange(h, device=x.device).view(1, h, 1)
            jj = torch.arange(w, device=x.device).view(1, 1, w)
            r0 = (d["ox"] - ch // 2).clamp(0, h - 1)
            r1 = (d["ox"] - ch // 2 + ch - 1).clamp(0, h - 1)
            c0 = (d["oy"] - cw // 2).clamp(0, w - 1)
            c1 = (d["oy"] - cw // 2 + cw - 1).clamp(0, w - 1)
            hole = (ii >= r0) & (ii <= r1) & (jj >= c0) & (jj <= c1)
            x = x * (~hole).to(x.dtype).unsqueeze(1)
    return x
