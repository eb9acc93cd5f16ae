"""BaseTrainer (trainers/base_trainer.py:12-131): model construction / weight init or load, iteration bookkeeping,
optimizer and LR-scheduler factories.  'adam' (the DefectGAN default, betas 0.5/0.999) is FlatAdam on the fused CUDA
kernel; under torchrun (WORLD_SIZE > 1) every network additionally gets a BucketReducer (overlapped NCCL all-reduce)."""
import math
from collections import defaultdict

import numpy as np
import torch
from torch import optim

from .. import functional as Fn
from .. import parallel
from ..models import create_model
from ..optim import FlatAdam, FlatRMSprop


class BaseTrainer:
    def __init__(self, opt):
        self.opt = opt
        self.rank, self.local_rank, self.world_size = parallel.init_distributed()
        self.model = create_model(opt)
        if getattr(opt, "continue_training", False):
            self.model.load("latest")
        elif getattr(opt, "load_model_name", None) is not None:
            self.model.load(opt.which_epoch)
        else:
            self.model.init_weights()
        for net in self.model.networks.values():  # identical replicas (no-op single process)
            parallel.broadcast_module(net)

        self.losses = defaultdict(list)
        self.dis_outputs = defaultdict(list)
        if opt.phase == "val":
            self.metrics = dict()

        self.iter_record_path = opt.ckpt_dir / opt.name / "iter.txt"
        self.first_epoch, self.iters = 1, 0
        assert hasattr(opt, "iters_per_epoch"), "opt must have attribute {iters_per_epoch}, it can be calculated by length of loader"
        if getattr(opt, "continue_training", False):
            self.first_epoch, self.iters = np.loadtxt(self.iter_record_path, delimiter=",", dtype=int)
        if opt.num_epochs == -1:
            opt.num_epochs = math.ceil(opt.num_iters / (opt.iters_per_epoch + 1e-12))
        opt.num_iters = opt.num_epochs * opt.iters_per_epoch
        assert self.first_epoch < opt.num_epochs, f"first_epoch {self.first_epoch} should not larger than num_epochs {opt.num_epochs}"
        assert self.iters < opt.num_iters, f"iters {self.iters} should not larger than num_iters {opt.num_iters}"
        opt.first_epoch = self.first_epoch

        self._init_lr(opt)
        self._create_optimizer(opt)
        self._create_scheduler(opt)

    def _init_lr(self, opt):
        self.lr = opt.lr[0]

    def _create_optimizer(self, opt):
        assert isinstance(self.lr, (int, float, dict)), "type of lr should be scalar or dict"
        if opt.optimizer == "adam":
            make = lambda ps, lr: FlatAdam(ps, lr=lr, betas=(0.5, 0.999))  # noqa: E731  base_trainer.py:79-81
        elif opt.optimizer == "rmsprop":
            make = lambda ps, lr: FlatRMSprop(ps, lr=lr)  # noqa: E731
        elif opt.optimizer == "adamw":  # base_trainer.py:82-84: optim.AdamW(betas=(0.9, 0.95)), torch's default weight_decay 1e-2
            make = lambda ps, lr: FlatAdam(ps, lr=lr, betas=(0.9, 0.95), weight_decay=1e-2, decoupled=True)  # noqa: E731
        elif opt.optimizer == "sgd":
            raise NotImplementedError("optimizer 'sgd' has no CUDA kernel on this path (adam | adamw | rmsprop)")
        else:
            raise NameError(f"optimizer named {opt.optimizer} not defined")
        self.optimizers, self.reducers = {}, {}
        for name, net in self.model.networks.items():
            lr = self.lr[name] if isinstance(self.lr, dict) else self.lr
            self.optimizers[name] = make(list(net.parameters()), lr)
            if self.world_size > 1:
                o = self.optimizers[name]
                self.reducers[name] = parallel.BucketReducer(o.flat_grads[0], o.slices(0))

    def _create_scheduler(self, opt):
        """base_trainer.py:91-126: 'step' (4 steps to lr_decay), 'exp', 'cos' (eta_min = lr * lr_decay); schedulers are
        advanced ``first_epoch`` times so that a resumed run continues at the right learning rate."""
        self.schedulers = {}
        for name, optimizer in self.optimizers.items():
            base_lr = self.lr[name] if isinstance(self.lr, dict) else self.lr
            if opt.scheduler == "step":
                sched = optim.lr_scheduler.StepLR(optimizer, step_size=opt.num_epochs // 4, gamma=opt.lr_decay ** (1 / 4))
            elif opt.scheduler == "exp":
                sched = optim.lr_scheduler.ExponentialLR(optimizer, gamma=opt.lr_decay ** (1 / opt.num_epochs))
            elif opt.scheduler == "cos":
                sched = optim.lr_scheduler.CosineAnnealingLR(optimizer, T_max=opt.num_epochs, eta_min=base_lr * opt.lr_decay)
            else:
                raise NameError(f"scheduler named {opt.scheduler} not defined")
            self.schedulers[name] = sched
        for _ in range(self.first_epoch):
            for sched in self.schedulers.values():
                sched.step()

    def _update_per_epoch(self, epoch=None):
        for sched in self.schedulers.values():
            sched.step()

    def _backward(self, loss):
        """loss.backward() with the weight gradients accumulated by the wgrad kernels themselves (Fn.accumulate_into_grad); the
        bucketed all-reduce overlap needs autograd's post-accumulate hooks and keeps the ordinary path."""
        with Fn.accumulate_into_grad(not (self.world_size > 1 and parallel.OVERLAP)):
            loss.backward()

    # ------------------------------------------------------------------------------------------------ data parallel
    def _reduce_and_step(self, name):
        """Join the overlapped bucket all-reduce of network ``name`` (if data parallel), then the fused optimizer step."""
        if name in self.reducers:
            self.reducers[name].finish()
        self.optimizers[name].step()
