"""DefectGanTrainer (trainers/defectgan_trainer.py:19-188): the D-step / G-step of the hot path and the epoch loop.

``_train_discriminator_once`` / ``_train_generator_once`` keep the reference's order of operations (zero_grad -> loss
graph -> weighted sum -> backward -> optimizer step -> ``.item()`` bookkeeping).  Differences, all invisible to the
results: the optimizer is FlatAdam (fused kernel), data-parallel gradient buckets are all-reduced under backward, and
during the G-step the discriminator's parameters do not require grad - the reference computes those gradients and
throws them away (``optimizers['D'].zero_grad()`` precedes D's next use, :171), we skip the wgrad kernels instead.
TensorBoard logging is optional (only if ``opt.log_dir`` is set and tensorboard is importable); the FID / IS / LPIPS
validation of ``--phase val`` needs downloaded weights and is out of the hot-path scope.
"""
import os
from collections import defaultdict

import numpy as np
import torch

from .base_trainer import BaseTrainer


class GraphedStep:
    """One trainer step (zero_grad .. optimizer step) as a CUDA graph per input-shape signature.

    Call 1 with a new signature runs the body eagerly (a real training step; it also triggers every lazy
    initialisation: kernel attributes, workspaces, RNG state).  Call 2 captures the body into a graph - capture
    executes nothing - and replays it; later calls copy the inputs into the static buffers, refresh the optimizer's
    device-side scalars (FlatAdam.advance) and replay.  Everything inside the body is device work on the capturing
    stream: libdei2i.so never synchronises or allocates, the noise maps come from torch's graph-safe Philox state,
    spectral-norm u/v, BN running stats and Adam state are updated in place at fixed addresses."""

    def __init__(self, body, optimizer, enabled, pool, staged=None, reuse_staged=False):
        self.body, self.optimizer, self.enabled, self.pool = body, optimizer, enabled, pool
        self.reuse_staged = reuse_staged  # only the G-step reuses (the D-step of every iteration reads the host batch afresh)
        self.states = {}
        # host tensors already on the device: (data_ptr, version, shape, dtype) -> static buffer of the step that copied them last.
        # The D-step and the G-step of one iteration are fed the same batch (defectgan_trainer.py:91-112): the second step takes
        # a device-to-device copy instead of a second trip over PCIe.  Shared between the two GraphedSteps of a trainer.
        self.staged = staged if staged is not None else {}
        self._copy_stream, self._landing, self._prefetched = None, {}, None

    @staticmethod
    def _key(t):
        return (t.data_ptr(), t._version, tuple(t.shape), t.dtype)

    def prefetch(self, *inputs):
        """Start the host-to-device copy of the NEXT call's inputs on a side stream, so that it runs under the step in flight instead
        of in front of its own (8 ranks copying 50 MB each at the same instant share PCIe switches; with a gradient all-reduce per step
        the slowest copy of every step is exposed on every rank).  The call that later receives the same pinned tensors takes a
        device-to-device copy from the landing buffers.  Pageable tensors are left to the ordinary path."""
        if not self.enabled or not all((not t.is_cuda) and t.is_pinned() for t in inputs):
            return
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream()
        sig = tuple((tuple(t.shape), t.dtype) for t in inputs)
        bufs = self._landing.get(sig)
        if bufs is None:
            dev = (self.optimizer[0] if isinstance(self.optimizer, (list, tuple)) else self.optimizer).flat_params[0].device
            bufs = self._landing[sig] = [torch.empty(t.shape, dtype=t.dtype, device=dev) for t in inputs]
        # the landing buffers may still be the source of the previous call's device-to-device copies on the compute stream
        self._copy_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._copy_stream):
            for b, t in zip(bufs, inputs):
                b.copy_(t, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        self._prefetched = ({self._key(t): b for t, b in zip(inputs, bufs)}, ev)

    def _stage(self, statics, inputs):
        fresh = {}
        landed, ev = self._prefetched if self._prefetched is not None else ({}, None)
        waited = False
        for s, t in zip(statics, inputs):
            key = self._key(t)
            src = self.staged.get(key) if (self.reuse_staged and not t.is_cuda) else None
            if src is None and not t.is_cuda and key in landed:
                if not waited:
                    torch.cuda.current_stream().wait_event(ev)
                    waited = True
                src = landed[key]
            s.copy_(src if src is not None else t, non_blocking=True)
            if not t.is_cuda:
                fresh[key] = s
        if waited:
            self._prefetched = None
        self.staged.clear()
        self.staged.update(fresh)

    def __call__(self, *inputs):
        if not self.enabled:
            return self.body(*inputs)
        key = tuple((tuple(t.shape), t.dtype) for t in inputs)
        st = self.states.get(key)
        if st is None:
            self.states[key] = {}
            return self.body(*inputs)
        if "graph" not in st:
            first = self.optimizer[0] if isinstance(self.optimizer, (list, tuple)) else self.optimizer
            dev = first.flat_params[0].device
            st["static"] = [torch.empty(t.shape, dtype=t.dtype, device=dev) for t in inputs]
            self._stage(st["static"], inputs)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, pool=self.pool[0]):
                st["out"] = self.body(*st["static"])
            if self.pool[0] is None:
                self.pool[0] = graph.pool()
            st["graph"] = graph
        else:
            self._stage(st["static"], inputs)
        for o in (self.optimizer if isinstance(self.optimizer, (list, tuple)) else (self.optimizer,)):
            o.advance()
        st["graph"].replay()
        return st["out"]


class DefectGanTrainer(BaseTrainer):
    def __init__(self, opt):
        super().__init__(opt)
        assert len(opt.loss_weight) == 5, f"length of loss weights must be 5, not {len(opt.loss_weight)}"
        self.loss_weights = dict(zip(("clf_d", "clf_g", "rec", "sd_cyc", "sd_con"), opt.loss_weight))
        self.loss_types = ["gan", "clf", "aux"]
        if opt.style_norm_block_type == "sean" and opt.style_distill:
            self.loss_types.append("distill")  # never reached: DefectGanGenerator refuses style_distill (normalization.SEAN)
        self._init_losses()
        if opt.phase == "val":
            raise NotImplementedError("--phase val (FID / IS / LPIPS with downloaded Inception weights) is out of the hot-path scope")
        self._d_params = list(self.model.netD.parameters())
        # whole-step CUDA graphs: ~1200 launches per step become one graph launch
        use_graph = getattr(opt, "cuda_graph", None)
        if use_graph is None:
            use_graph = os.environ.get("DEI2I_CUDA_GRAPH", "1") != "0"
        use_graph = bool(use_graph) and torch.device(opt.device).type == "cuda"
        # the style vector of the AdaIN variant comes from a third network, netE, stepped with the generator
        # (defectgan_trainer.py:140-141,162-163)
        self._has_e = "E" in self.optimizers
        # host-side draws inside the loss graph cannot be captured: SEAN picks stored embeddings with python's `random.choices`
        # (defectgan_model.py:408), the sean_alpha == 0 StyleExtractor draws its noise on the host (extractor.py:85)
        kind = opt.style_norm_block_type
        if (kind == "sean" and opt.sean_alpha != 0) or (kind == "adain" and opt.sean_alpha == 0):
            use_graph = False
        pool, staged = [None], {}
        g_opts = [self.optimizers["G"], self.optimizers["E"]] if self._has_e else self.optimizers["G"]
        self._graphed = {"D": GraphedStep(self._d_body, self.optimizers["D"], use_graph, pool, staged),
                         "G": GraphedStep(self._g_body, g_opts, use_graph, pool, staged, reuse_staged=True)}

    def _init_lr(self, opt):
        assert len(opt.lr) in (1, 2), f"length of lr must be 1 or 2, not {len(opt.lr)}"
        self.lr = {"D": opt.lr[0], "G": opt.lr[1]} if len(opt.lr) == 2 else opt.lr[0]
        if isinstance(self.lr, dict) and opt.style_norm_block_type == "adain":
            # the reference's dict has no 'E' entry and _create_optimizer raises KeyError for netE (base_trainer.py:86-88); the
            # extractor is stepped with the generator, so it takes the generator's rate (mae_trainer.py:35 does the same)
            self.lr["E"] = opt.lr[1]

    def _init_losses(self):
        self.losses = {t: defaultdict(list) for t in self.loss_types}

    # ------------------------------------------------------------------------------------------------ the hot path
    def _g_body(self, bg_data, df_labels, df_data):
        """zero_grad -> loss graph -> weighted sum -> backward -> (all-reduce) -> Adam; returns the 5 loss scalars stacked."""
        self.optimizers["G"].zero_grad()
        if self._has_e:
            self.optimizers["E"].zero_grad()
        for p in self._d_params:
            p.requires_grad_(False)
        try:
            gan, clf, rec, cyc, con = self.model("generator", bg_data, df_labels, df_data)
            w = self.loss_weights
            g_loss = gan + clf * w["clf_g"] + rec * w["rec"] + cyc * w["sd_cyc"] + con * w["sd_con"]
            self._backward(g_loss)
        finally:
            for p in self._d_params:
                p.requires_grad_(True)
        self._reduce_and_step("G")
        if self._has_e:
            self._reduce_and_step("E")
        return torch.stack([gan.detach(), clf.detach(), rec.detach(), cyc.detach(), con.detach()])

    def _d_body(self, bg_data, df_labels, df_data):
        self.optimizers["D"].zero_grad()
        gan, clf = self.model("discriminator", bg_data, df_labels, df_data)
        d_loss = gan + clf * self.loss_weights["clf_d"]
        self._backward(d_loss)
        self._reduce_and_step("D")
        return torch.stack([gan.detach(), clf.detach()])

    def _train_generator_once(self, bg_data, df_labels, df_data):
        """defectgan_trainer.py:138-168."""
        vals = self._graphed["G"](bg_data, df_labels, df_data).tolist()  # one D2H sync
        self.losses["gan"]["G"].append(vals[0])
        self.losses["clf"]["G"].append(vals[1])
        self.losses["aux"]["rec"].append(vals[2])
        self.losses["aux"]["cyc"].append(vals[3])
        self.losses["aux"]["con"].append(vals[4])

    def _train_discriminator_once(self, bg_data, df_labels, df_data, upcoming=None):
        """defectgan_trainer.py:170-180.  ``upcoming``: the next iteration's (bg_data, df_labels, df_data), if the caller already holds it -
        its host-to-device copy is started under this step (GraphedStep.prefetch)."""
        out = self._graphed["D"](bg_data, df_labels, df_data)
        if upcoming is not None:
            self._graphed["D"].prefetch(*upcoming)
        vals = out.tolist()
        self.losses["gan"]["D"].append(vals[0])
        self.losses["clf"]["D"].append(vals[1])

    # ------------------------------------------------------------------------------------------------ epoch loop
    def train(self, train_loaders, val_loaders=None):
        """defectgan_trainer.py:75-89; epochs run from first_epoch to num_epochs inclusive."""
        writer = self._make_writer()
        for epoch in range(self.first_epoch, self.opt.num_epochs + 1):
            self._init_losses()
            self._train_epoch(train_loaders, epoch)
            if writer is not None:
                self._write_tf_log(writer, epoch)
            if epoch % self.opt.save_ckpt_freq == 0 and self.rank == 0:
                self.model.save(epoch)
            self._update_per_epoch(epoch)
        if writer is not None:
            writer.close()

    def _train_epoch(self, data_loaders, epoch):
        """defectgan_trainer.py:91-120 (progress bar omitted): one D-step per iteration, one G-step every
        ``num_critics`` iterations, 'latest' checkpoint + iter.txt every ``save_latest_freq`` iterations."""
        def batches():
            for df_data, df_labels, _ in data_loaders["defects"]:
                bg_data, _, _ = next(data_loaders["background"])
                yield bg_data[: df_data.size(0)], df_labels, df_data

        it = batches()
        cur = next(it, None)
        while cur is not None:
            nxt = next(it, None)  # one batch of look-ahead: its copy to the device runs under this iteration's steps
            self.iters += 1
            bg_data, df_labels, df_data = cur
            self._train_discriminator_once(bg_data, df_labels, df_data, upcoming=nxt)
            if self.iters % self.opt.num_critics == 0:
                self._train_generator_once(bg_data, df_labels, df_data)
            if self.iters % self.opt.save_latest_freq == 0 and self.rank == 0:
                self.model.save("latest")
                np.savetxt(self.iter_record_path, (epoch, self.iters), fmt="%i", delimiter=",")
            cur = nxt

    def _make_writer(self):
        if self.rank != 0 or getattr(self.opt, "log_dir", None) is None:
            return None
        try:
            from torch.utils.tensorboard import SummaryWriter
        except Exception:
            return None
        return SummaryWriter(self.opt.log_dir / self.opt.name)

    def _write_tf_log(self, writer, epoch):
        for t in self.loss_types:
            writer.add_scalars(f"Losses/{t}", {k: sum(v) / (len(v) + 1e-12) for k, v in self.losses[t].items()}, epoch)
        writer.add_scalars("Lr", {f"net_{n}": s.get_last_lr()[0] for n, s in self.schedulers.items()}, epoch)

    def _update_per_epoch(self, epoch=None):
        super()._update_per_epoch(epoch)
        self.model.update_per_epoch(epoch)
