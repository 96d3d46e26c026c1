#!/usr/bin/env python
"""bench.py -- the hot path of htem/raygun on B200: one step = one pass of the 3-D generator
(forward + backward + loss) over one batch of synthetic EM-like uint8 volumes.

Workload at N=1 (BASELINE.json configs[1], "cfg2"): G-ResNet9 = Sequential(ResNet(3, input_nc=1, ngf=64,
n_blocks=9, InstanceNorm3d, padding 'same'), Conv3d(64,1,1), Tanh) forward+backward on a 2x1x128^3 patch,
bf16 tensor-core path with fp32 accumulation / fp32 master weights, SmoothL1 loss against a synthetic target.
N>1: the same per-GPU batch on every rank (weak scaling), gradients all-reduced with NCCL before they would
be consumed by the optimizer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2|cfg2-small]

Prints ONE JSON line (rank 0).  `value` is device-timed with inputs resident in HBM; `e2e` goes through the
public API from pinned HOST uint8 buffers (H2D + u8 normalise + step + D2H of the loss) every step.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOADS = {
    # name: (batch, side, n_blocks, ngf)
    "cfg2": (2, 128, 9, 64),
    "cfg2-small": (1, 64, 9, 64),
    # cfg5 (BASELINE configs[4]): deep G-UNet, 4 down-sampling levels, fmap_inc_factor 3, ngf 64 (1.78 B parameters),
    # forward + backward on one 1x1x256^3 patch; n_blocks = 0 selects the U-Net branch of run_ours
    "cfg5": (1, 256, 0, 64),
}
DEEP_KW = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=3, num_downs=5, padding_type="same",
               final_activation="Tanh")


def conv_flops_resnet(batch, side, n_blocks, ngf, n_down=2):
    """Algorithmic FLOPs (2 x MACs of Conv3d / ConvTranspose3d) of G-ResNet forward; SURVEY 8(d) convention."""
    fl = 0.0
    s = side
    fl += 2.0 * batch * s ** 3 * ngf * 1 * 343                    # stem 7^3 1->ngf
    c = ngf
    for _ in range(n_down):
        s //= 2
        fl += 2.0 * batch * s ** 3 * (2 * c) * c * 27
        c *= 2
    fl += n_blocks * 2 * (2.0 * batch * s ** 3 * c * c * 27)
    for _ in range(n_down):
        fl += 2.0 * batch * s ** 3 * c * (c // 2) * 27               # ConvT k3 s2: MACs = in_voxels*cin*cout*27
        s *= 2
        c //= 2
    fl += 2.0 * batch * s ** 3 * ngf * ngf * 343                    # head 7^3 ngf->ngf
    fl += 2.0 * batch * s ** 3 * ngf * 1                           # 1x1x1 -> 1
    return fl


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(tf_burst=p["bf16_tflops"], tf_sustained=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    hbm=p["hbm_gbs"], source="measured")
    return dict(tf_burst=1590.0, tf_sustained=1400.0, hbm=6650.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=mx, samples=len(sm), reasons=sorted(reasons))


def synth_batch(batch, side, seed):
    from raygun_b200.synth import em_like_u8
    import numpy as np
    return torch.from_numpy(np.stack([em_like_u8((side, side, side), seed=seed + i)[None] for i in range(batch)]))


# ---------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle port (the reference's algorithm restated over plain
# torch.nn.functional, oracle/restate.py) on the host cores
# ---------------------------------------------------------------------------------------------
def deep_unet_cpu_fn(side=32):
    """cfg5 on the host cores: the oracle's G-UNet forward + SmoothL1 + all gradients on one 1x1x32^3 patch."""
    from oracle import restate as O
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(42)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("unet", dict(DEEP_KW))
    sd = {k: v.detach().requires_grad_(v.is_floating_point()) for k, v in g.state_dict().items()}
    spec = O.unet_spec(input_nc=1, ngf=64, fmap_inc_factor=3, downsample_factors=[(2, 2, 2)] * 4, padding_type="same")
    x = O.u8_normalise(synth_batch(1, side, 42))
    tgt = O.u8_normalise(synth_batch(1, side, 43))
    leaves = [v for v in sd.values() if v.requires_grad]

    def step():
        y = O.generator_forward(sd, "unet", spec, x, final_activation="Tanh")
        loss = O.smooth_l1(tgt, y)
        torch.autograd.grad(loss, leaves, allow_unused=True)
        return float(loss.detach())

    return step, side ** 3


def cpu_step_fn(n_blocks, ngf, side, batch=1):
    if n_blocks == 0:
        return deep_unet_cpu_fn(32)
    from oracle import restate as O
    import raygun_b200.networks as N  # constructors only (parameter shapes / init); arithmetic is the oracle's
    torch.manual_seed(42)
    g = torch.nn.Sequential(N.ResNet(3, input_nc=1, ngf=ngf, n_blocks=n_blocks, padding_type="same"),
                            torch.nn.Conv3d(ngf, 1, 1), torch.nn.Tanh())
    with contextlib.redirect_stdout(sys.stderr):      # init_weights prints, as the reference's does
        N.init_weights(g, "kaiming")
    sd = {k: v.detach().clone().requires_grad_(v.is_floating_point()) for k, v in g.state_dict().items()}
    spec = O.resnet_spec(input_nc=1, ngf=ngf, n_blocks=n_blocks, padding_type="same",
                         norm=dict(kind="instance", affine=False))
    x = O.u8_normalise(synth_batch(batch, side, 42))
    tgt = O.u8_normalise(synth_batch(batch, side, 43))
    leaves = [v for v in sd.values() if v.requires_grad]

    def step():
        y = O.generator_forward(sd, "resnet", spec, x, final_activation="Tanh")
        loss = O.smooth_l1(tgt, y)
        torch.autograd.grad(loss, leaves, allow_unused=True)
        return float(loss.detach())

    return step, batch * side ** 3


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if args.workload == "cfg3":
        step, vox = cyclegan_cpu_fn(32)
        step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()
        dt = (time.perf_counter() - t0) / args.steps
        val = 1.0 / (dt * 8 * (64 ** 3 / vox))
        sample = "one oracle CycleStep on one 1x32^3 sample pair per step, scaled x64 to the global batch of 8 x 64^3"
        print(json.dumps(dict(metric="cyclegan_train_steps_per_sec", value=val, unit="steps/s", impl="reference",
                              n_gpus=args.gpus, steps=args.steps, warmup=1, ms_per_step=1e3 / val, higher_is_better=True,
                              scaling="strong", vs_baseline=None, dtype="f32", data="synthetic",
                              config=dict(workload="cfg3: CycleGAN step, CPU oracle port", sample=sample),
                              cpu_baseline=dict(value=val, unit="steps/s", cores=cores, kind="port", sample=sample),
                              e2e=dict(value=val, unit="steps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))))
        return
    if args.workload == "cfg4":
        step, vox = unet_cpu_fn(64)
        step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()
        dt = (time.perf_counter() - t0) / args.steps
        sample = "cfg1: G-UNet forward on one 1x1x64^3 patch per step, fp32 (fully convolutional; full workload is blockwise 256^3 reads)"
        print(json.dumps(dict(metric="generator_inference_voxels_per_sec", value=vox / dt, unit="voxels/s", impl="reference",
                              n_gpus=args.gpus, steps=args.steps, warmup=1, ms_per_step=dt * 1e3, higher_is_better=True,
                              scaling="strong", vs_baseline=None, dtype="f32", data="synthetic",
                              config=dict(workload="cfg4: blockwise G-UNet prediction, CPU oracle port", sample=sample),
                              cpu_baseline=dict(value=vox / dt, unit="voxels/s", cores=cores, kind="port", sample=sample),
                              e2e=dict(value=vox / dt, unit="voxels/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))))
        return
    batch, side, n_blocks, ngf = WORKLOADS[args.workload]
    sample_side = 32 if n_blocks == 0 else (48 if side >= 128 else 32)
    step, vox = cpu_step_fn(n_blocks, ngf, sample_side)
    for _ in range(max(1, min(args.warmup, 1))):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = vox / dt
    sample = f"1x1x{sample_side}^3 patch per step (network is fully convolutional; full workload {batch}x1x{side}^3)"
    out = dict(metric="generator_train_voxels_per_sec", value=val, unit="voxels/s", impl="reference", n_gpus=args.gpus,
               steps=args.steps, warmup=args.warmup, ms_per_step=dt * 1e3, higher_is_better=True, scaling="weak",
               vs_baseline=None, dtype="f32", data="synthetic",
               config=dict(workload=f"{args.workload}: {'deep G-UNet' if n_blocks == 0 else f'G-ResNet{n_blocks}'} "
                                    "fwd+bwd+SmoothL1, CPU oracle port", sample=sample),
               cpu_baseline=dict(value=val, unit="voxels/s", cores=cores, kind="port", sample=sample),
               e2e=dict(value=val, unit="voxels/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(out))


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist

    import raygun_b200 as rb
    from raygun_b200 import _C
    from raygun_b200 import conv as CV
    from raygun_b200 import functional as F
    from raygun_b200 import kernels as K
    from raygun_b200.cyclegan import make_generator

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (raygun_b200 has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rb.set_precision("bf16")
    batch, side, n_blocks, ngf = WORKLOADS[args.workload]

    torch.manual_seed(42)
    with contextlib.redirect_stdout(sys.stderr):
        if n_blocks == 0:
            g = make_generator("unet", dict(DEEP_KW)).to(dev)
            args.no_graph = True        # one 450 ms step of ~280 launches: launch gaps are noise, and the graph's
                                        # private pool would double the 58 GB activation footprint
        else:
            g = make_generator("resnet", dict(input_nc=1, output_nc=1, ngf=ngf, n_blocks=n_blocks, padding_type="same",
                                              final_activation="Tanh")).to(dev)
    params = [p for p in g.parameters()]
    host_x = synth_batch(batch, side, 42 + 10 * rank).pin_memory()
    host_t = synth_batch(batch, side, 43 + 10 * rank).pin_memory()
    x_dev = K.u8_normalise(host_x.to(dev), dtype=torch.float32)
    t_dev = K.u8_normalise(host_t.to(dev), dtype=torch.float32)

    def fwd_bwd(x, t):
        for p in params:
            p.grad = None
        y = g(x)
        loss = F.smooth_l1_loss(t, y)
        loss.backward()
        return loss

    def sync_grads():
        if world > 1:   # data-parallel: average gradients over ranks (flat bucket, NCCL over NVLink)
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat.div_(world)

    def step(x, t):
        loss = fwd_bwd(x, t)
        sync_grads()
        return loss

    # The step is ~300 kernel launches of 5 us - 10 ms each: capture normalise + forward + backward + loss in ONE CUDA
    # graph (static uint8 input buffers) so launch gaps do not sit between the kernels.  --no-graph keeps eager launches.
    graph = None
    xu_s = host_x.to(dev)
    tu_s = host_t.to(dev)
    if not args.no_graph:
        try:
            for _ in range(3):
                fwd_bwd(K.u8_normalise(xu_s, dtype=torch.float32), K.u8_normalise(tu_s, dtype=torch.float32))
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                loss_s = fwd_bwd(K.u8_normalise(xu_s, dtype=torch.float32), K.u8_normalise(tu_s, dtype=torch.float32))
            graph.replay()
            torch.cuda.synchronize()
        except Exception as exc:      # noqa: BLE001 -- fall back to eager launches, and say so in the JSON line
            print(f"bench.py: CUDA graph capture failed ({exc!r}); running eager", file=sys.stderr)
            graph = None
            torch.cuda.synchronize()

    def graph_step():
        graph.replay()
        sync_grads()
        return loss_s

    def e2e_step():
        if graph is not None:
            xu_s.copy_(host_x, non_blocking=True)
            tu_s.copy_(host_t, non_blocking=True)
            loss = graph_step()
        else:
            xu = host_x.to(dev, non_blocking=True)
            tu = host_t.to(dev, non_blocking=True)
            loss = step(K.u8_normalise(xu, dtype=torch.float32), K.u8_normalise(tu, dtype=torch.float32))
        return float(loss.detach())     # D2H of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms

    for _ in range(max(args.warmup, 3)):
        step(x_dev, t_dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = _C.launch_count
    step(x_dev, t_dev)                  # count the launches of one (eager) step; the graph replays the same kernels
    launches = _C.launch_count - l0 + (2 if graph is not None else 0)       # + the two u8_normalise launches in the graph
    ms = timed(graph_step if graph is not None else (lambda: step(x_dev, t_dev)), args.steps)
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e = timed(e2e_step, args.steps)

    # roofline of the dominant kernel family (tcgen05 implicit-GEMM convolutions), timed per launch with
    # CUDA events on the launching stream over one extra step
    CV.profile = []
    step(x_dev, t_dev)
    torch.cuda.synchronize()
    prof, CV.profile = CV.profile, None
    by_kind, by_kernel = {}, {}
    for kind, fl, a, b, kname in prof:
        dt = a.elapsed_time(b)
        for table, key in ((by_kind, kind), (by_kernel, kname)):
            d = table.setdefault(key, [0.0, 0.0, 0])
            d[0] += fl
            d[1] += dt
            d[2] += 1
    peaks = load_peaks()
    tot_fl = sum(v[0] for v in by_kind.values())
    tot_ms = sum(v[1] for v in by_kind.values())
    # the dominant kernel = the one with the largest share of the step; achieved = its algorithmic FLOPs / its time
    dom = max(by_kernel, key=lambda k: by_kernel[k][1])
    dfl, dms, dn = by_kernel[dom]
    achieved = dfl / (dms * 1e-3) / 1e12 if dms > 0 else 0.0
    # each launch is timed between its own pair of events in an instrumented (eager) step: the kernel runs at the
    # clocks of a kernel timed alone, so the denominator is the BURST peak (against the sustained figure the ring
    # kernel reads 1.01); frac_of_sustained is kept beside it
    peak = peaks["tf_burst"]
    traffic = None
    try:        # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture (profiles/)
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            traffic = json.load(f).get(dom, {}).get("dram_bytes_per_launch")
    except OSError:
        pass
    roofline = dict(bound="tensor", kernel=f"{dom} (tcgen05; {dn} launches per step)", achieved=achieved, peak=peak,
                    unit="TFLOP/s", frac=achieved / peak, frac_of_sustained=achieved / peaks["tf_sustained"],
                    peak_kind=f"{peaks['source']} burst bf16 (sustained {peaks['tf_sustained']})", traffic=traffic,
                    algorithmic_tflop_per_launch=dfl / dn / 1e12, ms_per_launch=dms / dn, share_of_step=dms / ms,
                    all_convs=dict(tflops=tot_fl / (tot_ms * 1e-3) / 1e12 if tot_ms > 0 else 0.0,
                                   frac=(tot_fl / (tot_ms * 1e-3) / 1e12) / peak if tot_ms > 0 else 0.0,
                                   launches=sum(v[2] for v in by_kind.values()), share_of_step=tot_ms / ms),
                    by_kernel={k: dict(tflops=v[0] / (v[1] * 1e-3) / 1e12, ms=v[1], launches=v[2]) for k, v in by_kernel.items()},
                    by_kind={k: dict(tflops=v[0] / (v[1] * 1e-3) / 1e12, ms=v[1], launches=v[2]) for k, v in by_kind.items()})

    if rank == 0:
        vox = batch * side ** 3 * world
        # cfg5: the algorithmic FLOPs of the step = sum over the conv launches (each counts 2 x MACs of its layer)
        fwd = conv_flops_resnet(batch, side, n_blocks, ngf) if n_blocks else tot_fl / 3
        net_name = f"G-ResNet{n_blocks} (ngf {ngf}, InstanceNorm, reflect 'same')" if n_blocks else \
            "deep G-UNet (4 down-sampling levels, fmap_inc_factor 3, ngf 64, 1.78 B parameters, reflect 'same')"
        csample = 32 if n_blocks == 0 else (48 if side >= 128 else 32)
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        cstep, cvox = cpu_step_fn(n_blocks, ngf, csample)
        cstep()
        t0 = time.perf_counter()
        cstep()
        cdt = time.perf_counter() - t0
        out = dict(metric="generator_train_voxels_per_sec", value=vox / (ms * 1e-3), unit="voxels/s", n_gpus=world,
                   steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=ms, higher_is_better=True, scaling="weak",
                   vs_baseline=None, dtype="bf16", data="synthetic",
                   config=dict(workload=f"{args.workload}: {net_name} fwd+bwd+SmoothL1 on {batch}x1x{side}^3 per GPU",
                               peak_hbm_gb=torch.cuda.max_memory_allocated() / 1e9,
                               l2="activation working set (GBs) >> 126 MB L2; no explicit flush",
                               parallelism=f"dp{world}" if world > 1 else "single", algorithmic_tflop_per_step=3 * fwd / 1e12,
                               cuda_graph=graph is not None),
                   step_tflops=3 * fwd * world / (ms * 1e-3) / 1e12,
                   e2e=dict(value=vox / (ms_e2e * 1e-3), unit="voxels/s", ms_per_step=ms_e2e,
                            h2d_bytes_per_step=int(host_x.numel() + host_t.numel()), d2h_bytes_per_step=4),
                   gpu_launches=int(launches), roofline=roofline,
                   cpu_baseline=dict(value=cvox / cdt, unit="voxels/s", cores=cores, kind="port",
                                     sample=f"1x1x{csample}^3 fwd+bwd, oracle/restate.py, 1 warm-up + 1 timed"),
                   clocks=clocks, conv_paths=dict(CV.stats))
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# cfg4: blockwise G-UNet prediction (generator inference voxels/s), blocks sharded over ranks, no collective
# ---------------------------------------------------------------------------------------------
UNET_FLOP_PER_VOXEL = 1.850e6          # G-UNet forward, algorithmic (halo recompute is overhead, not credit)


def unet_cpu_fn(side=64):
    """cfg1: the reference-runnable anchor -- G-UNet forward on a 1x1x64^3 patch, fp32, host cores (oracle port)."""
    from oracle import restate as O
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(42)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3,
                                        padding_type="same", final_activation="Tanh"))
    sd = {k: v.detach() for k, v in g.state_dict().items()}
    spec = O.unet_spec(input_nc=1, ngf=64, fmap_inc_factor=2, downsample_factors=[(2, 2, 2)] * 2, padding_type="same")
    x = O.u8_normalise(synth_batch(1, side, 42))

    def step():
        with torch.no_grad():
            return O.generator_forward(sd, "unet", spec, x, final_activation="Tanh")

    return step, side ** 3


def run_predict(args):
    import torch.distributed as dist

    import raygun_b200 as rb
    from raygun_b200 import _C
    from raygun_b200 import conv as CV
    from raygun_b200 import predict as P
    from raygun_b200.cyclegan import make_generator
    from raygun_b200.synth import em_like_u8
    import numpy as np

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (raygun_b200 has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rb.set_precision("bf16")
    R = int(args.read)                                 # read block; write block W = R - 48 (halo 24 >= 22, origins = 0 mod 4:
    W = R - 48                                         # SURVEY 8c); recompute overhead (R/W)^3 = 1.34 at R = 512 (88 GiB), 1.30 at R = 576 (125 GiB)
    nb = (4, 4, 2)                                     # write blocks per axis: 32 blocks, a 880 x 880 x 464 volume
    shape = tuple(n * W + (R - W) for n in nb)
    torch.manual_seed(42)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3,
                                        padding_type="same", final_activation="Tanh")).to(dev).eval()
    if R % 16 or R < 96:
        raise SystemExit("--read must be a multiple of 16, >= 96")
    tile = em_like_u8((232, 220, 220), seed=42)        # synthetic EM-like texture, mirrored-tiled to the volume size
    reps = [-(-s // t) for s, t in zip(shape, tile.shape)]
    vol_np = np.tile(np.concatenate([tile, tile[::-1]], 0), (reps[0], 1, 1))[: shape[0]]
    vol_np = np.tile(np.concatenate([vol_np, vol_np[:, ::-1]], 1), (1, reps[1], 1))[:, : shape[1]]
    vol_np = np.tile(np.concatenate([vol_np, vol_np[:, :, ::-1]], 2), (1, 1, reps[2]))[:, :, : shape[2]]
    vol = torch.from_numpy(np.ascontiguousarray(vol_np)).pin_memory()
    out = torch.zeros_like(vol).pin_memory()
    vol_dev = vol.to(dev)
    out_dev = torch.zeros_like(vol_dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            r = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms, r

    resident = lambda: P.predict_volume(g, vol_dev, (R,) * 3, (W,) * 3, out=out_dev, rank=rank, world=world, dtype=torch.bfloat16)
    through_host = lambda: P.predict_volume(g, vol, (R,) * 3, (W,) * 3, out=out, rank=rank, world=world, dtype=torch.bfloat16)
    for _ in range(max(1, min(args.warmup, 3)) if rank >= 0 else 0):
        resident()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = _C.launch_count
    ms, (_, nblk, vox) = timed(resident, args.steps)
    launches = (_C.launch_count - l0) // args.steps
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, _ = timed(through_host, args.steps)
    tot = torch.tensor([float(vox), float(nblk)], device=dev)
    if world > 1:
        dist.all_reduce(tot)
    vox_all, nblk_all = float(tot[0]), int(tot[1])
    if rank == 0:
        peaks = load_peaks()
        alg = vox_all * UNET_FLOP_PER_VOXEL / (ms * 1e-3) / 1e12
        exe = nblk_all * R ** 3 * UNET_FLOP_PER_VOXEL / (ms * 1e-3) / 1e12
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        cstep, cvox = unet_cpu_fn(64)
        cstep()
        t0 = time.perf_counter()
        cstep()
        cdt = time.perf_counter() - t0
        out_line = dict(metric="generator_inference_voxels_per_sec", value=vox_all / (ms * 1e-3), unit="voxels/s", n_gpus=world,
                        steps=args.steps, warmup=max(1, min(args.warmup, 3)), ms_per_step=ms, higher_is_better=True,
                        scaling="strong", vs_baseline=None, dtype="bf16", data="synthetic",
                        config=dict(workload=f"cfg4: blockwise G-UNet prediction, read {R}^3 / write {W}^3 (halo 24), "
                                             f"{nblk_all} blocks of a {shape[0]}x{shape[1]}x{shape[2]} uint8 volume per step",
                                    l2="every block's activations (GBs) >> 126 MB L2; no explicit flush",
                                    parallelism=f"blocks round-robin over {world} ranks, no collective"),
                        e2e=dict(value=vox_all / (ms_e2e * 1e-3), unit="voxels/s", ms_per_step=ms_e2e,
                                 h2d_bytes_per_step=int(nblk_all * R ** 3), d2h_bytes_per_step=int(vox_all)),
                        gpu_launches=int(launches),
                        roofline=dict(bound="tensor", kernel="whole block pipeline (ring_conv_kernel dominant)", achieved=alg,
                                      executed_incl_halo=exe, peak=peaks["tf_sustained"] * world, unit="TFLOP/s",
                                      frac=alg / (peaks["tf_sustained"] * world), frac_executed=exe / (peaks["tf_sustained"] * world),
                                      peak_kind=f"{peaks['source']} sustained bf16 x {world} GPUs (burst {peaks['tf_burst']})", traffic=None),
                        cpu_baseline=dict(value=cvox / cdt, unit="voxels/s", cores=cores, kind="port",
                                          sample="cfg1: G-UNet forward on 1x1x64^3, fp32, oracle/restate.py, 1 warm-up + 1 timed"),
                        clocks=clocks, conv_paths=dict(CV.stats))
        print(json.dumps(out_line))
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# cfg3: full CycleGAN training step (2 x G-UNet + 2 x NLayerDiscriminator, LSGAN + SmoothL1 cycle loss, Adam),
# global batch 8 of 1x64^3 patches per domain, data-parallel over batch with one gradient all-reduce per optimizer
# ---------------------------------------------------------------------------------------------
CYCLE_FLOP_PER_SAMPLE = 5.92e12        # 12 G-UNet passes + 16 D passes, algorithmic
G_KW = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same", final_activation="Tanh")
D_KW = dict(input_nc=1, ngf=64, n_layers=3, kw=3, downsampling_kw=2)


def cyclegan_cpu_fn(side=32):
    """One oracle CycleStep (reference LinkCycleLoss semantics, oracle/restate.py) on one sample pair, host cores."""
    import functools
    from oracle import restate as O
    import raygun_b200.networks as N
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(42)
    IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    with contextlib.redirect_stdout(sys.stderr):
        gs = [make_generator("unet", dict(G_KW)) for _ in range(2)]
        ds = [N.NLayerDiscriminator(3, norm_layer=IN, **D_KW) for _ in range(2)]
        for d in ds:
            N.init_weights(d, "kaiming")
    sds = [{k: v.detach().clone() for k, v in m.state_dict().items()} for m in gs + ds]
    uspec = O.unet_spec(input_nc=1, ngf=64, fmap_inc_factor=2, downsample_factors=[(2, 2, 2)] * 2, padding_type="same")
    dspec = O.nld_spec(input_nc=1, ngf=64, n_layers=3, norm=dict(kind="instance", affine=False), kw=3, downsampling_kw=2)
    cs = O.CycleStep(*sds, g_fwd=lambda sd, x: O.generator_forward(sd, "unet", uspec, x, final_activation="Tanh"),
                     d_fwd=lambda sd, x: O.nld3d_forward(sd, dspec, x))
    a = O.u8_normalise(synth_batch(1, side, 42))
    b = O.u8_normalise(synth_batch(1, side, 43))
    return (lambda: cs.step(a, b)), side ** 3


def run_cyclegan(args):
    import functools
    import torch.distributed as dist

    import raygun_b200 as rb
    from raygun_b200 import _C
    from raygun_b200 import conv as CV
    from raygun_b200 import kernels as K
    from raygun_b200 import networks as N
    from raygun_b200.cyclegan import CycleGANStep, make_generator
    from raygun_b200.dist import GradSync

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (raygun_b200 has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rb.set_precision("bf16")
    GLOBAL_BATCH, SIDE = 8, 64
    if GLOBAL_BATCH % world:
        raise SystemExit("cfg3: the global batch of 8 must divide over the ranks")
    per = GLOBAL_BATCH // world
    torch.manual_seed(42)                       # identical initial weights on every rank
    IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    with contextlib.redirect_stdout(sys.stderr):
        g1, g2 = (make_generator("unet", dict(G_KW)).to(dev) for _ in range(2))
        d1, d2 = (N.NLayerDiscriminator(3, norm_layer=IN, **D_KW).to(dev) for _ in range(2))
        for d in (d1, d2):
            N.init_weights(d, "kaiming")
    sync = GradSync({"G": list(g1.parameters()) + list(g2.parameters()),
                     "D": list(d1.parameters()) + list(d2.parameters())}) if world > 1 else None
    # the whole step (forwards, backwards, gradient all-reduces, both Adam updates) replays as ONE CUDA graph after
    # three eager steps; --no-graph keeps ~1500 eager launches per step (host-bound at one sample per GPU)
    step = CycleGANStep(g1, g2, d1, d2, lr=1e-5, grad_sync=sync, graph=not args.no_graph)
    host_a = synth_batch(per, SIDE, 42 + 100 * rank).pin_memory()
    host_b = synth_batch(per, SIDE, 43 + 100 * rank).pin_memory()
    dev_a = K.u8_normalise(host_a.to(dev), dtype=torch.bfloat16)
    dev_b = K.u8_normalise(host_b.to(dev), dtype=torch.bfloat16)

    def resident():
        return step.step(dev_a, dev_b)

    def e2e():
        a = K.u8_normalise(host_a.to(dev, non_blocking=True), dtype=torch.bfloat16)
        b = K.u8_normalise(host_b.to(dev, non_blocking=True), dtype=torch.bfloat16)
        out = step.step(a, b)
        return float(out["G1"])                 # D2H of a loss every step, as the reference's logging does

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            r = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms, r

    for _ in range(max(3, args.warmup) + (2 if not args.no_graph else 0)):   # 3 eager steps, the capture, replays
        resident()
    if step.graph_error is not None:
        print(f"bench.py: CUDA graph capture failed ({step.graph_error}); running eager", file=sys.stderr)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = _C.launch_count
    ms, losses = timed(resident, args.steps)
    launches = step.graph_launches if step.graph is not None else (_C.launch_count - l0) // args.steps
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, _ = timed(e2e, args.steps)
    if rank == 0:
        peaks = load_peaks()
        alg = GLOBAL_BATCH * CYCLE_FLOP_PER_SAMPLE / (ms * 1e-3) / 1e12
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        cstep, cvox = cyclegan_cpu_fn(32)
        cstep()
        t0 = time.perf_counter()
        cstep()
        cdt = time.perf_counter() - t0
        cpu_steps = 1.0 / (cdt * GLOBAL_BATCH * (SIDE ** 3 / cvox))      # scaled to 8 samples of 64^3 (fully convolutional)
        line = dict(metric="cyclegan_train_steps_per_sec", value=1e3 / ms, unit="steps/s", n_gpus=world, steps=args.steps,
                    warmup=max(3, args.warmup), ms_per_step=ms, higher_is_better=True, scaling="strong", vs_baseline=None,
                    dtype="bf16", data="synthetic",
                    config=dict(workload="cfg3: CycleGAN step, 2 x G-UNet + 2 x NLayerDiscriminator(n_layers 3, kw 3), LSGAN + "
                                         "SmoothL1 cycle loss (lambda 10/0/1), Adam lr 1e-5, global batch 8 of 1x64^3 per domain",
                                per_gpu_batch=per, l2="activations of one step (GBs) >> 126 MB L2; no explicit flush",
                                parallelism=f"dp{world}: one gradient all-reduce per optimizer step (G1+G2, D1+D2)",
                                cuda_graph=step.graph is not None),
                    e2e=dict(value=1e3 / ms_e2e, unit="steps/s", ms_per_step=ms_e2e,
                             h2d_bytes_per_step=int(2 * per * SIDE ** 3), d2h_bytes_per_step=4),
                    gpu_launches=int(launches),
                    roofline=dict(bound="tensor", kernel="whole step (ring_conv_kernel / halo_wgrad_kernel dominant)", achieved=alg,
                                  peak=peaks["tf_sustained"] * world, unit="TFLOP/s", frac=alg / (peaks["tf_sustained"] * world),
                                  peak_kind=f"{peaks['source']} sustained bf16 x {world} GPUs (burst {peaks['tf_burst']})", traffic=None,
                                  algorithmic_tflop_per_step=GLOBAL_BATCH * CYCLE_FLOP_PER_SAMPLE / 1e12),
                    cpu_baseline=dict(value=cpu_steps, unit="steps/s", cores=cores, kind="port",
                                      sample="one oracle CycleStep on one 1x32^3 sample pair, scaled x64 to 8 samples of 64^3"),
                    clocks=clocks, losses={k: float(v) for k, v in losses.items()}, conv_paths=dict(CV.stats))
        print(json.dumps(line), flush=True)
    if world > 1:
        # a CUDA graph that holds captured NCCL collectives must be gone before its communicator is destroyed
        # (ncclCommDestroy waits for it otherwise); a timer guards the exit path of every rank all the same
        import gc
        import threading
        threading.Timer(90.0, lambda: os._exit(0)).start()
        step.graph = None
        del step
        gc.collect()
        barrier()
        dist.destroy_process_group()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS) + ["cfg3", "cfg4"])
    ap.add_argument("--read", type=int, default=512, help="cfg4: read block size (multiple of 16; write block = read - 48)")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying one CUDA graph")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "cfg4":
        run_predict(args)
    elif args.workload == "cfg3":
        run_cyclegan(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
