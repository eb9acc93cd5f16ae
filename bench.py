#!/usr/bin/env python
"""Benchmark of the DefectGAN hot path: one "step" = one _train_discriminator_once + one _train_generator_once
(num_critics = 1) on a synthetic batch; metric = images / second (B * n_gpus / step time).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench.py --gpus N ...

Ours: DefectGanTrainer (de_i2i_gan_b200.trainers) -> mirrored modules -> libdei2i.so (sm_100a kernels), BF16 storage /
tcgen05 MMA with FP32 accumulation, workload BASELINE.json configs[1]: 256x256, batch 32 per GPU.
  value : device-timed (CUDA events) over K steps with the inputs already resident in HBM;
  e2e   : the same K steps through the same trainer API fed from PINNED HOST tensors (H2D of both image batches and
          the labels inside the timed region, losses read back to the host every step);
  roofline : every convolution launch of the timed region is bracketed by CUDA events on the launching stream;
          achieved = sum(algorithmic FLOPs) / sum(durations) of the dominant kernel class vs MEASURED_PEAKS.json;
  cpu_baseline : the CPU restatement of the reference (oracle/, plain torch FP32) timed on this box's host cores on a
          bounded sample (one D+G step at 256x256, batch 2).
Reference arm (--impl reference): the same CPU restatement timed for K steps with all host threads (the reference
itself is Python under /root/reference, which does not exist on the GPU box).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "DefectGAN train img/s at 256x256 (D-step + G-step, batch 32/GPU)"
UNIT = "img/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="per-GPU batch (BASELINE.json configs[1]: 32)")
    ap.add_argument("--image-size", type=int, default=256)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-batch", type=int, default=2)
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs only: skip the pinned-host pass")
    ap.add_argument("--spade-dense", action="store_true",
                    help="evaluate SPADE gamma|beta with the dense full-resolution convolutions even for 1x1 label maps "
                         "(default: exact 5x5 border-class map, see DESIGN.md 4a)")
    ap.add_argument("--breakdown", default=None, help="write a per-convolution-geometry timing table (JSON) to this path")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(power), "samples": len(sm),
                "reasons": sorted(reasons)}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], burst=p["bf16_tflops"], sustained=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    source="MEASURED_PEAKS.json")
    return dict(hbm=6650.0, burst=1590.0, sustained=1400.0, source="fallback (B200_PROFILING.md)")


# ---------------------------------------------------------------------------------------------------- CPU arm
def cpu_oracle_steps(image_size, batch, steps, warmup, threads):
    """Times D-step + G-step of the CPU restatement (oracle/defectgan_ref.py, full-width nets, FP32).  Returns the
    list of per-step seconds."""
    import torch
    from oracle import defectgan_ref as R
    torch.set_num_threads(threads)
    g = R.random_generator_state(noise_weight=0.0)
    d = R.random_discriminator_state(image_size=image_size)
    tr = R.Trainer(g, d)
    bg, labels, df = R.synthetic_batch(batch, image_size, seed=0)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        tr.train_discriminator_once(bg, labels, df)
        tr.train_generator_once(bg, labels, df)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return times


def run_reference(args):
    """--impl reference: the reference's CPU path (restated in oracle/, see module docstring) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    bs = args.cpu_sample_batch
    # bound the whole run to a few minutes: probe one step, shrink the sample if needed
    probe = cpu_oracle_steps(args.image_size, bs, 1, 0, cores)[0]
    budget = 240.0
    while bs > 1 and probe * (args.steps + args.warmup) > budget:
        bs //= 2
        probe = cpu_oracle_steps(args.image_size, bs, 1, 0, cores)[0]
    steps, warmup = args.steps, args.warmup
    if probe * (steps + warmup) > budget:
        warmup = 0
        steps = max(1, min(steps, int(budget / probe)))
    times = cpu_oracle_steps(args.image_size, bs, steps, warmup, cores)
    sec = sum(times) / len(times)
    val = bs / sec
    sample = f"D-step + G-step at {args.image_size}x{args.image_size}, batch {bs} per step, FP32, torch CPU ({torch.__version__})"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"DefectGAN D+G training step {args.image_size}x{args.image_size}", "sample_batch": bs,
                   "device": "host CPU"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def _finish(world):
    """Multi-rank exit: every rank meets at a last barrier, flushes and leaves without the NCCL / CUDA-graph teardown
    (a rank that returns early while another sits in destroy_process_group() hangs the launcher)."""
    if world <= 1:
        return
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


# ---------------------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from de_i2i_gan_b200 import ops, options, parallel
    from de_i2i_gan_b200.trainers import find_trainer_using_model_name
    from oracle import defectgan_ref as R  # synthetic_batch only (seeded input factory); never on the compute path

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback (use --impl reference for the CPU arm)")
    if args.spade_dense:
        from de_i2i_gan_b200.models.networks import normalization
        normalization.CONSTANT_LABEL_CLASSES = False
    rank, local, world = parallel.init_distributed()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    B, H = args.batch, args.image_size

    torch.manual_seed(123)  # fix_rand_seed(123), train_defectgan.py:120 - identical initial weights on every rank
    opt = options.readme_recipe(image_size=H, batch_size=B, num_critics=1, num_epochs=10, iters_per_epoch=10 ** 6,
                                name="bench", ckpt_dir=__import__("pathlib").Path("/tmp/dei2i_bench_ckpt"))
    opt.device = dev
    trainer = find_trainer_using_model_name(opt.model)(opt)
    torch.manual_seed(1000 + rank)  # per-rank noise stream
    bg, labels, df = R.synthetic_batch(B, H, seed=rank)  # per-rank shard of the global batch
    host = [t.pin_memory() for t in (bg, labels, df)]
    resident = [t.to(dev) for t in host]
    # one trip over PCIe per iteration, as the reference's loop does (data.to(device) once, then D-step and G-step,
    # defectgan_trainer.py:91-112): the D-step copies bg, labels, df from the pinned host batch, the G-step of the same
    # iteration takes a device-to-device copy of them (GraphedStep._stage)
    h2d = sum(t.numel() * t.element_size() for t in host)
    d2h = 7 * 4                                               # 2 + 5 loss scalars

    def step(inputs):
        b, l, d = inputs
        trainer._train_discriminator_once(b, l, d)
        trainer._train_generator_once(b, l, d)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(inputs, steps, profile):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.PROFILE = [] if profile else None
        barrier()
        ev0.record()
        for _ in range(steps):
            step(inputs)
        ev1.record()
        barrier()
        prof, ops.PROFILE = ops.PROFILE, None
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms, prof

    for _ in range(max(args.warmup, 3)):  # call 1 eager, call 2 captures the D-step / G-step CUDA graphs, call 3+ replay
        step(resident)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms, _ = timed(resident, args.steps, False)
    clk = clocks.stop() if rank == 0 else None
    if args.skip_e2e:
        ms_e2e = float("nan")
    else:
        step(host)  # warm the pinned-host path once
        ms_e2e, _ = timed(host, args.steps, False)
    mem_gb = torch.cuda.max_memory_allocated() / 2 ** 30
    # per-kernel pass: a replayed CUDA graph cannot be bracketed kernel by kernel, so the SAME kernels are launched eagerly for
    # the same K steps right after the timed region, every convolution between two CUDA events on the launching stream
    graphed = any(g.enabled for g in trainer._graphed.values())
    for g in trainer._graphed.values():
        g.enabled = False
    step(resident)
    l0 = ops.launch_count()
    ms_eager, prof = timed(resident, args.steps, True)
    launches = ops.launch_count() - l0

    # ---- roofline of the convolution kernels from the in-band events
    agg = {}
    if args.breakdown and rank == 0:
        table = {}
        for kind, flops, e0, e1, tag in prof:
            t = table.setdefault(f"{kind} {tag}", [0, 0.0, 0])
            t[0] += flops; t[1] += e0.elapsed_time(e1); t[2] += 1
        rows = [dict(kernel=k, launches_per_step=v[2] / args.steps, ms_per_step=v[1] / args.steps, gflop_per_launch=v[0] / v[2] / 1e9,
                     tflops=(v[0] / (v[1] * 1e-3) / 1e12 if v[1] > 0 else 0.0)) for k, v in table.items()]
        rows.sort(key=lambda r: -r["ms_per_step"])
        with open(args.breakdown, "w") as f:
            json.dump(rows, f, indent=1)
    for kind, flops, e0, e1, _tag in prof:
        a = agg.setdefault(kind, [0, 0.0, 0])
        a[0] += flops
        a[1] += e0.elapsed_time(e1)
        a[2] += 1
    peaks = load_peaks()
    detail = {k: {"launches": v[2], "ms_per_step": v[1] / args.steps, "tflops": v[0] / (v[1] * 1e-3) / 1e12 if v[1] > 0 else 0.0,
                  "gflop_per_launch": v[0] / v[2] / 1e9} for k, v in agg.items()}
    conv_ms = sum(v[1] for v in agg.values())
    conv_flops = sum(v[0] for v in agg.values())
    # dominant kernel = the dense tcgen05 convolution class with the largest share of the step (tensor-bound); the skinny
    # classes (8-channel inputs / <= 4 outputs, HBM-bound by construction) and the tiny class (< 2 GFLOP launches: 5x5 SPADE
    # class-map convolutions, classifier heads - launch-latency bound) are listed in `detail`
    dense = {k: v for k, v in agg.items() if k.endswith("_dense")}
    dom = max(dense, key=lambda k: dense[k][1]) if dense else (max(agg, key=lambda k: agg[k][1]) if agg else None)
    roofline = None
    if dom is not None:
        a = detail[dom]
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tpath):  # dram__bytes_read.sum + dram__bytes_write.sum per launch from a committed `ncu --set full` capture
            with open(tpath) as f:
                traffic = json.load(f).get(dom)
        dense_ms = sum(v[1] for v in dense.values())
        dense_flops = sum(v[0] for v in dense.values())
        roofline = {"bound": "tensor", "kernel": dom + " (conv_tma_kernel / wgrad_tma_kernel, tcgen05)", "achieved": a["tflops"],
                    "peak": peaks["sustained"], "unit": "TFLOP/s", "frac": a["tflops"] / peaks["sustained"], "traffic": traffic,
                    "peak_source": peaks["source"] + " bf16_tflops_sustained (kernels timed inside a long step)",
                    "all_dense_conv_tflops": dense_flops / (dense_ms * 1e-3) / 1e12 if dense_ms > 0 else None,
                    "dense_conv_share_of_step": dense_ms / ms_eager,
                    "all_conv_tflops": conv_flops / (conv_ms * 1e-3) / 1e12, "conv_share_of_step": conv_ms / ms_eager,
                    "measured_in": "eager re-run of the same K steps directly after the timed region (identical kernels; a replayed CUDA graph cannot be bracketed per kernel)",
                    "eager_ms_per_step": ms_eager / args.steps, "detail": detail}
    if rank != 0:
        _finish(world)
        return
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        bs = args.cpu_sample_batch
        t = cpu_oracle_steps(H, bs, 1, 0, cores)[0]
        cpu = {"value": bs / t, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"1 D-step + G-step at {H}x{H}, batch {bs}, FP32 torch CPU restatement (oracle/defectgan_ref.py), {t:.1f} s"}
    sec = ms * 1e-3 / args.steps
    print(json.dumps({
        "metric": METRIC, "value": B * world / sec, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": f"DefectGAN full training step (1 D-step + 1 G-step) {H}x{H}, batch {B}/GPU, --add_noise --use_spectral",
                   "global_batch": B * world, "parallelism": f"dp{world}", "executed_conv_tflop_per_step_per_gpu": conv_flops / args.steps / 1e12,
                   "reference_dense_conv_tflop_per_step_per_gpu": 81.3147 * (B / 32) * (H / 256) ** 2,
                   "spade": ("dense full-resolution gamma|beta convolutions" if args.spade_dense else
                             "gamma|beta evaluated exactly on the 5x5 border-class map of the 1x1 training label map "
                             "(results identical; FLOPs of the dense evaluation are NOT credited - see executed vs reference TFLOP)"),
                   "l2": "no flush needed: every step streams its own activations (>> 126 MB L2)", "peak_mem_gb": round(mem_gb, 1),
                   "cuda_graph": graphed},
        "e2e": {"value": B * world / (ms_e2e * 1e-3 / args.steps), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches, "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
    }))
    _finish(world)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
