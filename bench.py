#!/usr/bin/env python
"""Headline benchmark: SwinIR-M gray-denoising TRAINING step (BASELINE.json configs[1], "cfg2") on N B200s.

    python bench.py --gpus N --steps K --warmup W                 # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K --warmup W # the reference arithmetic on the host CPU cores

One step = zero_grad -> forward -> Charbonnier loss -> backward -> Adam -> update_E (EMA 0.999), i.e.
`optimize_parameters` of the reference (models/model_plain.py:270-318 with options/swinir/train_swinir_denoising_gray.json)
on 8 patches of 1x128x128 per GPU.  `value` is patches/s over all ranks with inputs resident in HBM;
`e2e` is the same step fed from pinned host memory with the loss read back every step.  The roofline object
describes the kernel group that takes the largest share of the step, timed per launch with CUDA events on the
launching stream in an instrumented replay of the same launch lists right after the timed region.

/root/reference does not exist on the GPU box: the reference arm and the cpu_baseline leg time the oracle
(oracle/swinir_oracle.py, a functional restatement of the reference modules pinned against them by tests/).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CFG2 = dict(upscale=1, in_chans=1, img_size=128, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
            num_heads=[6] * 6, mlp_ratio=2, upsampler=None, resi_connection="1conv")
BATCH_PER_GPU = 8
PATCH = 128
FWD_GFLOP_PER_PATCH = 399.94          # SURVEY.md section 8(d), measured with hooks on the reference modules
REAL_DIMS = {192: 180, 576: 540, 384: 360, "head_dim": 30}
WORKLOAD = "cfg2: SwinIR-M gray-DN training (ws8, C180, 6x6 RSTB), SPECT-shaped 1x128x128 patches, batch 8/GPU"


# ------------------------------------------------------------------------------------------------ synthetic data

def spect_batch(batch, size, seed):
    """SPECT-shaped synthetic pair: lambda = smooth positive field of Gaussian blobs (mean ~5
    counts), H = Poisson(lambda)/150, L = 8*Binomial(Poisson(lambda), 1/8)/150 (linear /150 normalisation,
    utils/utils_spect.py:15-16 of the reference)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(size, dtype=torch.float32), torch.arange(size, dtype=torch.float32),
                            indexing="ij")
    lam = torch.full((batch, 1, size, size), 0.8)
    for b in range(batch):
        for _ in range(6):
            cy, cx = (torch.rand(2, generator=g) * size).tolist()
            sig = 6.0 + 20.0 * torch.rand(1, generator=g).item()
            amp = 4.0 + 36.0 * torch.rand(1, generator=g).item() ** 2
            lam[b, 0] += amp * torch.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * sig * sig))
    counts = torch.poisson(lam, generator=g)
    low = torch.binomial(counts, torch.full_like(counts, 0.125), generator=g) * 8.0
    return (low / 150.0).contiguous(), (counts / 150.0).contiguous()


def charbonnier(x, y, eps=1e-9):
    """models/loss.py:208-218 of the reference."""
    import torch
    d = x - y
    return torch.mean(torch.sqrt(d * d + eps))


# ------------------------------------------------------------------------------------------------ clocks

class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}

    def __init__(self, index, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.mask, self.max_mhz = [], 0, None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001 - reported in the JSON line
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    self.mask |= nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    self.mask |= nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join()
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "samples": len(s),
                "reasons": sorted(v for k, v in self.REASONS.items() if self.mask & k)}


def _physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except (ValueError, IndexError):
            return local_rank
    return local_rank


# ------------------------------------------------------------------------------------------------ oracle legs (CPU)

def oracle_train_steps(steps, warmup, batch, seed=0, budget_s=240.0):
    """The reference's training step restated on the oracle (CPU, fp32, all host threads).  Returns
    (seconds per step, steps actually timed, threads)."""
    import torch
    from oracle import swinir_oracle as so
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = so.SwinIRConfig(img_size=CFG2["img_size"], in_chans=1, embed_dim=180, depths=(6,) * 6, num_heads=(6,) * 6,
                          window_size=8, mlp_ratio=2, upscale=1, img_range=1.0, upsampler="")
    sd = so.deterministic_state(so.state_spec(cfg), seed, cfg)
    params = [v for k, v in sd.items() if v.dtype.is_floating_point and not k.endswith("attn_mask")]
    for p in params:
        p.requires_grad_(True)
    opt = torch.optim.Adam(params, lr=2e-4, betas=(0.9, 0.999), weight_decay=0)
    ema = [p.detach().clone() for p in params]       # netE; E_decay = 0.999 (train_swinir_denoising_gray.json:64)
    low, high = spect_batch(batch, PATCH, 1234)
    times = []
    t_begin = time.perf_counter()
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        out = so.swinir_forward(sd, low, cfg)
        loss = so.charbonnier(out, high)
        loss.backward()
        opt.step()
        with torch.no_grad():                        # update_E, models/model_base.py:247-252
            for e, p in zip(ema, params):
                e.mul_(0.999).add_(p.data, alpha=1 - 0.999)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
        if time.perf_counter() - t_begin > budget_s and times:
            break
    return sum(times) / len(times), len(times), threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_batch = 1
    sec, done, threads = oracle_train_steps(args.steps, args.warmup, sample_batch)
    value = sample_batch / sec
    sample = "%d patch(es) of the 8-patch batch per step, %d timed steps, fp32, oracle port of the reference " \
             "modules (reference tree is not on the GPU box)" % (sample_batch, done)
    line = {
        "impl": "reference", "metric": "SwinIR-M gray-DN train patches/s", "value": value, "unit": "patches/s",
        "n_gpus": args.gpus, "steps": done, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "patches/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "patches/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def _wrap_dp(model, local_rank, use_ddp):
    """Patch-batch data parallelism (models/model_base.py:113-119): torch DDP (--ddp) or the single flat all-reduce."""
    if use_ddp:
        from torch.nn.parallel import DistributedDataParallel as DDP
        return DDP(model, device_ids=[local_rank], broadcast_buffers=False, gradient_as_bucket_view=True)
    from kair_b200.parallel import FlatDataParallel
    return FlatDataParallel(model, overlap=not os.environ.get("KAIR_B200_NO_OVERLAP"))


def _dp_name(world, use_ddp):
    if world == 1:
        return "dp1"
    return "dp%d (%s)" % (world, "torch DDP" if use_ddp else "FlatDataParallel: NCCL all-reduce of the flat gradient "
                          "buffer, per-RSTB slices overlapped with the backward pass"
                          + (" [overlap off]" if os.environ.get("KAIR_B200_NO_OVERLAP") else ""))


def _ncu_traffic(tag):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of kernel group `tag`, from the committed
    `ncu --set full` capture (profiles/ncu_traffic.json, written by tools/ncu_summary.py); None if not captured."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        return d["bytes_per_launch"].get(tag)
    except Exception:  # noqa: BLE001
        return None


def _group_table(groups):
    """per kernel group: launches, total ms, algorithmic TFLOP/s, algorithmic GB/s and the written share of it"""
    return {k: {"n": v["count"], "ms": round(v["ms"], 4),
                "tflops": round(v["flops"] / (v["ms"] * 1e-3) / 1e12, 1) if v["ms"] > 0 else 0.0,
                "gbs": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 0) if v["ms"] > 0 else 0.0,
                "write_gbs": round(v.get("wbytes", 0.0) / (v["ms"] * 1e-3) / 1e9, 0) if v["ms"] > 0 else 0.0}
            for k, v in sorted(groups.items(), key=lambda kv: -kv[1]["ms"])}


# ------------------------------------------------------------------------------------------------ B200 arm

def run_b200(args):
    import torch
    import torch.distributed as dist
    from kair_b200 import profiling
    from kair_b200.network_swinir import SwinIR

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the sm_100a path has no CPU fallback "
                           "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    torch.manual_seed(0)
    model = SwinIR(**CFG2).to(dev).train()           # drop_path_rate: constructor default 0.1, as define_G leaves it
    if args.precision == "fp32":
        model.set_precision("fp32", args.terms)
    plan_args = (BATCH_PER_GPU, PATCH, PATCH, True, args.precision, args.terms)
    net = model
    if world > 1:
        net = _wrap_dp(model, local_rank, args.ddp)
    # optimize_parameters (models/model_plain.py:270-318) with E_decay = 0.999: loss, Adam, update_E.  Default: the fused
    # tail (kair_pixel_loss + kair_adam_ema, one launch each); --torch-tail: torch ops (fused Adam + foreach EMA).
    import copy
    from kair_b200.loss import CharbonnierLoss
    from kair_b200.optim import FusedAdamEMA, update_ema
    ema_params = [p.detach().clone() for p in model.parameters()]
    if args.torch_tail:
        opt = torch.optim.Adam(model.parameters(), lr=2e-4, betas=(0.9, 0.999), weight_decay=0, fused=True)
        lossfn = charbonnier
    else:
        opt = FusedAdamEMA(model.parameters(), lr=2e-4, betas=(0.9, 0.999), weight_decay=0, ema_params=ema_params,
                           ema_decay=0.999)
        lossfn = CharbonnierLoss(eps=1e-9)
    params_list = list(model.parameters())

    low_h, high_h = spect_batch(BATCH_PER_GPU, PATCH, 1234 + rank)
    low_h, high_h = low_h.pin_memory(), high_h.pin_memory()
    low, high = low_h.to(dev), high_h.to(dev)

    def step(lo, hi):
        opt.zero_grad(set_to_none=True)
        out = net(lo)
        loss = lossfn(out, hi)
        loss.backward()
        opt.step()
        if args.torch_tail:
            update_ema(ema_params, params_list, 0.999)
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    for _ in range(max(args.warmup, 3)):
        step(low, high)
    if args.ncu_slice:
        # profiling aid (never a bench value): replay ONE launch of every distinct kernel/shape group between
        # cudaProfilerStart/Stop so that `ncu --profile-from-start off --set full` captures ~30 launches, not 550
        import ctypes
        plan = model.engine(dev).plan(*plan_args)
        seen, pick = set(), []
        for l in list(plan.fwd) + list(plan.bwd):
            tag = profiling.describe(l, REAL_DIMS)[0]
            if tag not in seen:
                seen.add(tag)
                pick.append((tag, l))
        sth = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        for tag, l in pick:
            l.run(sth)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        print(json.dumps({"ncu_slice": [t for t, _ in pick]}), flush=True)
        return
    sampler = ClockSampler(_physical_gpu_index(local_rank))
    sampler.start()
    ms_total = timed(lambda: step(low, high), args.steps)
    clocks = sampler.stop()

    # end-to-end: inputs come from pinned host memory every step, the loss goes back to the host every step
    def e2e_step():
        low.copy_(low_h, non_blocking=True)
        high.copy_(high_h, non_blocking=True)
        return step(low, high).item()

    for _ in range(2):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)

    # host time per step: the CPU side of one step (Python + ctypes launch issue, no synchronisation inside) measured
    # from a drained device, so it is not hidden behind -- nor inflated by -- queued GPU work
    host_ms = []
    for _ in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step(low, high)
        host_ms.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
    host_ms_step = sorted(host_ms)[len(host_ms) // 2]

    patches = BATCH_PER_GPU * world
    ms_step = ms_total / args.steps
    value = patches / (ms_step * 1e-3)
    e2e_value = patches / (ms_e2e / args.steps * 1e-3)

    # ---- per-launch instrumented replay (rank 0): which kernel group dominates, and its achieved TFLOP/s
    plan = model.engine(dev).plan(*plan_args)
    launches = list(plan.fwd) + list(plan.bwd)
    n_pack = len(plan.Wt["pack"])
    roof, groups = None, None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        peak_tf = peaks.get("bf16_tflops_sustained")
        peak_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
        if not peak_tf:
            peak_tf, peak_src = 1400.0, "fallback (B200_PROFILING.md sustained ~1.4 PFLOP/s)"
        import ctypes
        sth = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        plan.gflat.zero_(); plan.dwflat.zero_()
        for t_ in plan._thin:
            t_.zero_()
        per = profiling.time_launches(launches, sth, reps=3)
        groups = profiling.summarize(launches, per, REAL_DIMS)
        tot = sum(d["ms"] for d in groups.values())
        top_tag, top = max(groups.items(), key=lambda kv: kv[1]["ms"])
        peak_gbs = peaks.get("hbm_gbs") or 6500.0
        roof = profiling.roofline_of(top, peak_tf, peak_gbs)
        roof.update({
                "traffic": _ncu_traffic(top_tag), "kernel": top_tag, "launches_per_step": top["count"],
                "avg_launch_ms": top["ms"] / top["count"], "share_of_kernel_time": top["ms"] / tot,
                "algorithmic_bytes_per_launch": top["bytes"] / top["count"],
                "algorithmic_flops_per_launch": top["flops"] / top["count"],
                "peak_source": peak_src + " / hbm_gbs",
                # pure-write ceiling of this GPU's HBM (fill kernel, measured live): write-heavy groups in `groups`
                # (write_gbs close to this number) are at their roofline although `gbs` is far from `peak`
                "hbm_write_peak": round(profiling.measure_fill_gbs(dev), 0),
                "whole_step": {"tflops": 3 * FWD_GFLOP_PER_PATCH * BATCH_PER_GPU / ms_step,
                               "frac": 3 * FWD_GFLOP_PER_PATCH * BATCH_PER_GPU / ms_step / peak_tf,
                               "sum_kernel_ms": tot},
                "groups": _group_table(groups)})

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sec, done, threads = oracle_train_steps(3, 1, 1, budget_s=60.0)
        cpu = {"value": 1.0 / sec, "unit": "patches/s", "cores": threads, "kind": "port",
               "sample": "1 patch (1x128x128) per step, %d timed steps after 1 warm-up, fp32 oracle port of the "
                         "reference modules incl. Adam and update_E" % done}

    # the inference half of BASELINE.json's metric ("... & DnCNN/SwinIR inference MPix/s") in the same process, about a
    # second each: cfg1 (B=1 CUDA-graph replay, and B=64) and cfg3, each with value, e2e, dominant-kernel roofline
    # fraction and its own CPU baseline (oracle port)
    n_launches = len(launches)
    also = None
    if rank == 0 and world == 1 and not args.no_also:
        del plan, launches
        also = {}
        for key, wl, b_, st_ in (("cfg1", "cfg1", 1, 200), ("cfg1_b64", "cfg1", 64, 20), ("cfg3", "cfg3", 32, 10)):
            sub = argparse.Namespace(**vars(args))
            sub.workload, sub.batch, sub.steps, sub.warmup, sub.precision = wl, b_, st_, 3, "bf16"
            sub.no_cpu_baseline = args.no_cpu_baseline or key == "cfg1_b64"
            try:
                l_ = run_b200_workload(sub, emit=False)
                also[key] = {"metric": l_["metric"], "value": l_["value"], "unit": l_["unit"],
                             "ms_per_step": l_["ms_per_step"], "steps": st_, "dtype": l_["dtype"],
                             "workload": l_["config"]["workload"], "weights": l_["config"].get("weights"),
                             "e2e": l_["e2e"], "gpu_launches": l_["gpu_launches"], "clocks": l_["clocks"],
                             "roofline": {k: v for k, v in l_["roofline"].items() if k != "groups"},
                             "cpu_baseline": l_["cpu_baseline"]}
            except Exception as e:  # noqa: BLE001 -- the headline line must not be lost to a side measurement
                also[key] = {"error": repr(e)}

    if rank == 0:
        line = {
            "metric": "SwinIR-M gray-DN train patches/s", "value": value, "unit": "patches/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else "bf16x%d split operands, fp32 accumulate (fp32-accurate mode)" % args.terms,
            "data": "synthetic", "host_ms_per_step": host_ms_step,
            "config": {"workload": WORKLOAD, "global_batch": patches, "parallelism": _dp_name(world, args.ddp),
                       "precision": args.precision + (" (bf16 tensor-core operands, fp32 accumulate; 2e-2 contract)" if args.precision == "bf16" else " (split operands, terms=%d; 1e-4 / 1e-3 contract)" % args.terms),
                       "drop_path_rate": 0.1, "optimizer": "Adam lr 2e-4 + EMA 0.999 (update_E)", "loss": "Charbonnier eps 1e-9",
                       "step_tail": "torch" if args.torch_tail else "fused (kair_pixel_loss, kair_adam_ema)",
                       "l2": "per-step working set (~13 GB of activations) far exceeds the 126 MB L2; no flush needed"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "patches/s",
                    "h2d_bytes_per_step": (low_h.numel() + high_h.numel()) * 4 * world,
                    "d2h_bytes_per_step": 4 * world, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": (n_launches + n_pack + (0 if args.torch_tail else 2)) * args.steps,
            "roofline": roof, "cpu_baseline": cpu, "also": also,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ other configs
# BASELINE.json configs[0], [2], [3], [4] (cfg1, cfg3, cfg4, cfg5 of SURVEY.md section 8(d)): `--workload cfgN`.
# The default line (no flag) stays cfg2, the config the headline metric is quoted on.

SWIN_SR = dict(upscale=4, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
               num_heads=[6] * 6, mlp_ratio=2, resi_connection="1conv")
WORKLOADS = {
    "cfg1": "cfg1: DnCNN-17 'R' gray sigma=25 inference (dncnn_25.pth architecture), synthetic 1x256x256 images, batch %d/GPU",
    "cfg3": "cfg3: SwinIR-M classical SR x4 (pixelshuffle) inference, synthetic 3x64x64 LR patches, batch %d/GPU",
    "cfg4": "cfg4: DRUNet (UNetRes 64-128-256-512, nb=4) training, 4-channel 128x128 patches (RGB + sigma map), batch %d/GPU, L1 + Adam 1e-4",
    "cfg5": "cfg5: SwinIR-M real-world SR x4 (nearest+conv) tiled inference, synthetic 3x2048x2048 image, tile 256 overlap 32 (100 tiles), %d tiles per forward, tiles sharded over ranks",
}
DEFAULT_BATCH = {"cfg1": 1, "cfg3": 32, "cfg4": 16, "cfg5": 4}


def _dncnn_weights(model):
    """dncnn_25.pth tensors from the committed known-answer fixture when present (data, not oracle code)."""
    import numpy as np
    import torch
    path = os.path.join(ROOT, "tests", "golden", "kat", "dncnn_25.npz")
    if os.path.exists(path):
        z = np.load(path)
        model.load_state_dict({str(k): torch.from_numpy(z["w:" + str(k)]) for k in z["keys"]}, strict=True)
        return "model_zoo/dncnn_25.pth (from tests/golden/kat)"
    return "random init"


def oracle_workload(name, batch, steps, warmup, budget_s=60.0):
    """CPU leg of a workload on the oracle port (fp32, all host threads), on a bounded sample; returns
    (units per second, timed steps, threads, sample description)."""
    import torch
    from oracle import convnets_oracle as co
    from oracle import swinir_oracle as so
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    if name == "cfg1":
        sd = so.deterministic_state(co.dncnn_state_spec(1, 1, 64, 17, "R"), 0)
        x = torch.rand(1, 1, 256, 256, generator=g)
        units, sample = 256 * 256 / 1e6, "1 image of 1x256x256 per step"
        def fn():
            with torch.no_grad():
                co.dncnn_forward(sd, x, nb=17, act_mode="R")
    elif name in ("cfg3", "cfg5"):
        up = "pixelshuffle" if name == "cfg3" else "nearest+conv"
        cfg = so.SwinIRConfig(img_size=64, in_chans=3, embed_dim=180, depths=(6,) * 6, num_heads=(6,) * 6, window_size=8,
                              mlp_ratio=2, upscale=4, img_range=1.0, upsampler=up)
        sd = so.deterministic_state(so.state_spec(cfg), 0, cfg)
        side = 64 if name == "cfg3" else 128
        x = torch.rand(1, 3, side, side, generator=g)
        units = side * side / 1e6
        sample = "1 LR patch of 3x%dx%d per step%s" % (side, side, "" if name == "cfg3" else
                                                      " (a quarter of one 256x256 tile; cost is linear in pixels)")
        def fn():
            with torch.no_grad():
                so.swinir_forward(sd, x, cfg)
    elif name == "cfg4":
        sd = so.deterministic_state(co.unetres_state_spec(4, 3, (64, 128, 256, 512), 4, False), 0)
        params = [v.requires_grad_(True) for v in sd.values()]
        opt = torch.optim.Adam(params, lr=1e-4)
        x, tgt = torch.rand(1, 4, 128, 128, generator=g), torch.rand(1, 3, 128, 128, generator=g)
        units, sample = 1.0, "1 patch of 4x128x128 per step (forward + backward + Adam)"
        def fn():
            opt.zero_grad(set_to_none=True)
            torch.mean(torch.abs(co.unetres_forward(sd, x, nb=4) - tgt)).backward()
            opt.step()
    else:
        raise SystemExit("unknown workload " + name)
    times, t_begin = [], time.perf_counter()
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        fn()
        if it >= warmup:
            times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_begin > budget_s and times:
            break
    sec = sum(times) / len(times)
    return units / sec, len(times), threads, sample + ", %d timed steps, fp32 oracle port of the reference modules" % len(times)


_W_METRIC = {"cfg1": ("DnCNN-17 inference MPix/s", "MPix/s"), "cfg3": ("SwinIR-M SR x4 inference LR-MPix/s", "MPix/s"),
             "cfg4": ("DRUNet train patches/s", "patches/s"), "cfg5": ("SwinIR-M SR x4 tiled inference LR-MPix/s", "MPix/s")}


def run_reference_workload(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    name = args.workload
    batch = args.batch or DEFAULT_BATCH[name]
    value, done, threads, sample = oracle_workload(name, batch, args.steps, min(args.warmup, 1), budget_s=200.0)
    metric, unit = _W_METRIC[name]
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus, "steps": done,
        "warmup": min(args.warmup, 1), "ms_per_step": None, "higher_is_better": True,
        "scaling": "strong" if name == "cfg5" else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOADS[name] % batch, "sample": sample},
        "cpu_baseline": {"value": value, "unit": unit, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
    }), flush=True)


def run_b200_workload(args, emit=True):
    import ctypes
    import torch
    import torch.distributed as dist
    from kair_b200 import profiling, tiling
    name = args.workload
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the sm_100a path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1 and emit:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    prec = getattr(args, "precision", "bf16")
    if prec == "fp32" and name not in ("cfg3", "cfg5"):
        raise SystemExit("--precision fp32 covers the SwinIR workloads (cfg2, cfg3, cfg5)")
    batch = args.batch or DEFAULT_BATCH[name]
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(1234 + rank)
    real_dims = dict(REAL_DIMS)
    weights = "random init (reference default initialisation, seed 0)"
    extra_cfg = {}
    n_tiles_rank = None

    if name == "cfg1":
        from kair_b200.network_dncnn import DnCNN
        model = DnCNN(1, 1, 64, 17, "R")
        weights = _dncnn_weights(model)
        model = model.to(dev).eval()
        for p_ in model.parameters():
            p_.requires_grad = False
        clean = torch.rand(batch, 1, 256, 256, generator=g)
        host = [(clean + torch.randn(clean.shape, generator=g) * (25.0 / 255.0)).pin_memory()]
        units = batch * 256 * 256 / 1e6
        plan_of = lambda: model.engine(dev).plan(batch, 256, 256, False, False)
        if args.eager:
            def step(x):
                with torch.no_grad():
                    return model(x)
        else:
            # launch-bound at batch 1 (17 kernels of ~16 us): replay the forward as one CUDA graph (kair_b200/graphs.py)
            from kair_b200.graphs import GraphedInference
            graphed = GraphedInference(model, host[0].to(dev))
            step = lambda x: graphed(x)
        fwd_gflop = 72.63 * batch
        real_dims = {}
    elif name == "cfg3":
        from kair_b200.network_swinir import SwinIR
        model = SwinIR(upsampler="pixelshuffle", **SWIN_SR).to(dev).eval().set_precision(prec, args.terms)
        host = [torch.rand(batch, 3, 64, 64, generator=g).pin_memory()]
        units = batch * 64 * 64 / 1e6
        plan_of = lambda: model.engine(dev).plan(batch, 64, 64, False, prec, args.terms)
        def step(x):
            with torch.no_grad():
                return model(x)
        fwd_gflop = 107.11 * batch
    elif name == "cfg4":
        from kair_b200.network_unet import UNetRes
        model = UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", downsample_mode="strideconv",
                        upsample_mode="convtranspose", bias=False).to(dev).train()
        net = model
        if world > 1:
            net = _wrap_dp(model, local_rank, args.ddp)
        # options/train_drunet.json: L1 loss, Adam 1e-4, no EMA (E_decay unset); fused tail unless --torch-tail
        if args.torch_tail:
            opt = torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
            l1 = lambda a, b: torch.mean(torch.abs(a - b))
        else:
            from kair_b200.loss import L1Loss
            from kair_b200.optim import FusedAdamEMA
            opt = FusedAdamEMA(model.parameters(), lr=1e-4)
            l1 = L1Loss()
        high = torch.rand(batch, 3, 128, 128, generator=g)
        sigma = torch.rand(batch, 1, 1, 1, generator=g) * 50.0 / 255.0
        low = torch.cat([high + torch.randn(high.shape, generator=g) * sigma, sigma.expand(batch, 1, 128, 128)], 1)
        host = [low.contiguous().pin_memory(), high.pin_memory()]
        units = float(batch)
        plan_of = lambda: model.engine(dev).plan(batch, 128, 128, True)
        def step(lo, hi):
            opt.zero_grad(set_to_none=True)
            loss = l1(net(lo), hi)
            loss.backward()
            opt.step()
            return loss
        fwd_gflop = 3 * 69.39 * batch
        real_dims = {}
    elif name == "cfg5":
        from kair_b200.network_swinir import SwinIR
        model = SwinIR(upsampler="nearest+conv", **SWIN_SR).to(dev).eval().set_precision(prec, args.terms)
        side = args.image or 2048
        img = tiling.pad_to_window(torch.rand(1, 3, side, side, generator=torch.Generator().manual_seed(1234)), 8)
        host = [img.pin_memory()]
        units = side * side / 1e6 / world          # whole job = one image; each rank does its share of the tiles
        n_tiles = len(tiling.tile_origins(img.shape[2], img.shape[3], 256, 32))
        n_tiles_rank = len([i for i in range(n_tiles) if i % world == rank])
        extra_cfg = {"tiles": n_tiles, "tiles_on_rank0": n_tiles_rank, "padded": list(img.shape[2:])}
        plan_of = lambda: model.engine(dev).plan(batch, 256, 256, False, prec, args.terms)
        def step(x):
            with torch.no_grad():
                E, Wt, _ = tiling.tiled_partial(model, x, 256, 32, 4, 8, tile_batch=batch, rank=rank, world=world)
                if world > 1:      # merge of the per-rank partial sums on rank 0 (the only exchange; after the compute)
                    dist.reduce(E, 0)
                    dist.reduce(Wt, 0)
                return E.div_(Wt) if rank == 0 else E
        fwd_gflop = 1791.1 * n_tiles_rank
    else:
        raise SystemExit("unknown workload " + name)

    devin = [h.to(dev) for h in host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        out = step(*devin)
    sampler = ClockSampler(_physical_gpu_index(local_rank))
    sampler.start()
    ms_total = timed(lambda: step(*devin), args.steps)
    clocks = sampler.stop()
    out_bytes = out.numel() * out.element_size() if out.dim() > 0 else 4
    host_out = torch.empty(out.shape, dtype=out.dtype).pin_memory() if out.dim() > 0 else None

    def e2e_step():
        for d_, h_ in zip(devin, host):
            d_.copy_(h_, non_blocking=True)
        o_ = step(*devin)
        if host_out is None:
            return o_.item()
        if name != "cfg5" or rank == 0:
            host_out.copy_(o_, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    ms_step = ms_total / args.steps
    value = units * world / (ms_step * 1e-3)
    e2e_value = units * world / (ms_e2e / args.steps * 1e-3)

    roof, cpu = None, None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        peak_tf = peaks.get("bf16_tflops_sustained") or 1400.0
        peak_gbs = peaks.get("hbm_gbs") or 6500.0
        plan = plan_of()
        launches = list(plan.fwd) + list(plan.bwd or [])
        if plan.bwd:
            plan.grads.flat.zero_()
            for t_ in plan.zero_before_bwd:
                t_.zero_()
        sth = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        per = profiling.time_launches(launches, sth, reps=3)
        groups = profiling.summarize(launches, per, real_dims)
        tot = sum(d["ms"] for d in groups.values())
        top_tag, top = max(groups.items(), key=lambda kv: kv[1]["ms"])
        roof = profiling.roofline_of(top, peak_tf, peak_gbs)
        passes = 1 if name != "cfg5" else -(-n_tiles_rank // batch)
        roof.update({"traffic": None, "kernel": top_tag, "launches_per_step": top["count"] * passes,
                     "avg_launch_ms": top["ms"] / top["count"], "share_of_kernel_time": top["ms"] / tot,
                     "algorithmic_bytes_per_launch": top["bytes"] / top["count"],
                     "algorithmic_flops_per_launch": top["flops"] / top["count"],
                     "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)",
                     "hbm_write_peak": round(profiling.measure_fill_gbs(dev), 0),
                     "whole_step": {"tflops": fwd_gflop / ms_step, "frac": fwd_gflop / ms_step / peak_tf,
                                    "sum_kernel_ms": tot * passes},
                     "groups": _group_table(groups)})
        n_launch = len(launches) * passes
        if world == 1 and not args.no_cpu_baseline:
            v, done, threads, sample = oracle_workload(name, batch, 3, 1, budget_s=40.0)
            cpu = {"value": v, "unit": _W_METRIC[name][1], "cores": threads, "kind": "port", "sample": sample}
        metric, unit = _W_METRIC[name]
        line = {
            "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if name == "cfg5" else "weak",
            "vs_baseline": None,
            "dtype": "bf16" if prec == "bf16" else "bf16x%d split operands, fp32 accumulate (fp32-accurate mode)" % args.terms,
            "data": "synthetic",
            "config": dict({"workload": WORKLOADS[name] % batch, "precision": prec, "parallelism": _dp_name(world, args.ddp) if name == "cfg4" else "replicas x%d, no collective" % world
                            if name != "cfg5" else "tiles sharded over %d ranks, one reduce of the partial sums" % world,
                            "weights": weights,
                            "l2": "inputs/activations of a step exceed the 126 MB L2 except cfg1 at batch 1 "
                                  "(whole net L2-resident by design); no flush"}, **extra_cfg),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": unit, "h2d_bytes_per_step": sum(h.numel() * 4 for h in host) * world,
                    "d2h_bytes_per_step": out_bytes * (1 if name == "cfg5" else world),
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": n_launch * args.steps, "roofline": roof, "cpu_baseline": cpu,
        }
        if not emit:
            return line
        print(json.dumps(line), flush=True)
    if world > 1 and emit:
        dist.destroy_process_group()
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="cfg2", choices=["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"],
                    help="BASELINE.json config to run; cfg2 (SwinIR training) is the headline and the default")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override for cfg1/3/4, tiles per forward for cfg5")
    ap.add_argument("--image", type=int, default=0, help="cfg5: image side (default 2048)")
    ap.add_argument("--ddp", action="store_true", help="N>1 training: wrap in torch DistributedDataParallel (the "
                    "reference's wrapper) instead of kair_b200.parallel.FlatDataParallel")
    ap.add_argument("--eager", action="store_true", help="cfg1: launch the forward kernel by kernel instead of "
                    "replaying the captured CUDA graph")
    ap.add_argument("--torch-tail", action="store_true", help="cfg2: torch loss / fused Adam / foreach EMA instead of "
                    "the fused kair_pixel_loss + kair_adam_ema tail")
    ap.add_argument("--ncu-slice", action="store_true", help="profiling aid: one launch per kernel group in a "
                    "cudaProfilerStart/Stop range")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"], help="bf16: tensor-core operands in bf16 "
                    "(default, 2e-2 contract); fp32: the fp32-accurate split-operand mode (1e-4 / 1e-3 contract) of the "
                    "SwinIR workloads")
    ap.add_argument("--terms", type=int, default=3, choices=[3, 6], help="--precision fp32: component products per GEMM")
    ap.add_argument("--no-also", action="store_true", help="cfg2: skip the cfg1 / cfg3 side measurements (`also`)")
    args = ap.parse_args()
    if args.workload != "cfg2":
        (run_reference_workload if args.impl == "reference" else run_b200_workload)(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
