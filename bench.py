#!/usr/bin/env python
"""bench.py -- monocular-depth init hot path (align + subsample + unproject) on N B200s.

A "step" is one pass of the hot path over one batch of synthetic views (BASELINE.json config 2:
Mip-NeRF-360-shaped, 200 views of 1297x840, 5000 SfM points per view + 20 % foreign, RANSAC
alignment + adaptive subsampling + default SfM-density patch mask).  N > 1 is weak scaling by
default: every rank owns its own 200 views (no data-path collective) and the exchange step -- every
rank ends up with the whole scene cloud in view order -- is fused into the emit kernel over
NVLink.  `--scaling strong` shards the workload's views over the N ranks instead (BASELINE configs
3 / 4: `--workload c3_... / c4_... --scaling strong`).

  python bench.py [--gpus N --steps K --warmup W]            our arm (CUDA kernels)
  python bench.py --impl reference [...]                     the reference's CPU path (oracle port)
  python bench.py --postprocess lof|merge [...]              C5 post-processing leg (a12-a14)
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
PKG = "3dgs_monocular_depth_init_b200"

WORKLOADS = {
    # name: (views per GPU, H, W, sfm points per view, aligner, subsample)
    "c2_mipnerf360_200x1297x840_ransac_adaptive": (200, 840, 1297, 5000, "ransac", "adaptive"),
    "c1_single_1297x840_lstsq_stride10": (1, 840, 1297, 5000, "lstsqrs", 10),
    "c3_tanks_300x1920x1080_ransac_adaptive": (300, 1080, 1920, 5000, "ransac", "adaptive"),
    "c4_4k_500x3840x2160_ransac_adaptive": (500, 2160, 3840, 5000, "ransac", "adaptive"),
    "c5_large_2000x1600x1200_ransac_adaptive": (2000, 1200, 1600, 5000, "ransac", "adaptive"),
}
DEFAULT_WORKLOAD = "c2_mipnerf360_200x1297x840_ransac_adaptive"


def mod(name=""):
    return importlib.import_module(PKG + ("." + name if name else ""))


def make_config(aligner, subsample):
    cfg = mod("config").Config()
    cfg.mdi.alignment.aligner = mod("depth_alignment.config").DepthAlignmentStrategyEnum(aligner)
    cfg.mdi.subsample_factor = subsample
    return cfg


class ClockSampler:
    """SM clock / throttle-reason samples during the timed region: NVML from a thread (a few
    microseconds per query; a looping nvidia-smi process was seen to disturb the PCIe-bound end-to-end
    steps), nvidia-smi only if NVML cannot be loaded."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None
        self.sm, self.mx, self.reasons, self._stop, self.thread, self.source = [], [], set(), False, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            bits = {"hw_slowdown": pynvml.nvmlClocksEventReasonHwSlowdown if hasattr(pynvml, "nvmlClocksEventReasonHwSlowdown")
                    else pynvml.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": getattr(pynvml, "nvmlClocksEventReasonHwThermalSlowdown",
                                                   getattr(pynvml, "nvmlClocksThrottleReasonHwThermalSlowdown", 0)),
                    "sw_thermal_slowdown": getattr(pynvml, "nvmlClocksEventReasonSwThermalSlowdown",
                                                   getattr(pynvml, "nvmlClocksThrottleReasonSwThermalSlowdown", 0)),
                    "sw_power_cap": getattr(pynvml, "nvmlClocksEventReasonSwPowerCap",
                                            getattr(pynvml, "nvmlClocksThrottleReasonSwPowerCap", 0))}
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons",
                                  getattr(pynvml, "nvmlDeviceGetCurrentClocksThrottleReasons", None))
            self.mx.append(float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)))

            def loop():
                while not self._stop:
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        if get_reasons is not None:
                            r = int(get_reasons(h))
                            for n, b in bits.items():
                                if b and (r & int(b)):
                                    self.reasons.add(n)
                    except Exception:  # noqa: BLE001
                        pass
                    time.sleep(0.02)

            self.thread = threading.Thread(target=loop, daemon=True)
            self.thread.start()
            self.source = "nvml"
            return
        except Exception:  # noqa: BLE001 - no NVML binding: fall back to the nvidia-smi loop
            self.thread = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
            self.source = "nvidia-smi"
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.thread is not None:
            self._stop = True
            self.thread.join(timeout=1.0)
            sm = sorted(self.sm)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                    "reasons": sorted(self.reasons), "samples": len(sm), "source": "nvml"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


def algorithmic_bytes(kernel: str, HW: int, V: int, n_kept: int, n_sfm_slots: int, adaptive: bool,
                      lattice_row_fraction: float = 0.6004) -> float:
    """Algorithmic bytes of one launch of a kernel, per SURVEY.md 8(d): per view
    B = P*HW*5 + HW + 36*Nk + 16*M, split over the kernels that move them.  lattice_row_fraction: share of
    the image rows that some stride of the default range 5..15 divides (counted: 0.6004)."""
    return {
        "stats_main_kernel": V * HW * 5.0,            # the one statistics read of depth + mask
        "keep_lattice_kernel": V * HW * 5.0,          # the mask pass's read of depth + mask (rows without
                                                      # a lattice point are skipped: real traffic is lower)
        "keep_rows_kernel": V * HW * 5.0 * lattice_row_fraction,   # depth + mask of the rows some stride divides
        "keep_zero_fill": V * HW * 1.0,               # keep mask written
        "emit_points_kernel": 36.0 * n_kept,          # rgb read, xyz + rgb written
        "project_filter_kernel": 16.0 * n_sfm_slots,  # sfm xyz + index
    }.get(kernel, 0.0)


def workload_config(workload: str, scaling: str, world: int) -> dict:
    """The `config` object of the JSON line: identical in both arms for the same command line."""
    V, H, W, n_sfm, aligner, subsample = WORKLOADS[workload]
    per_gpu = V if scaling == "weak" else -(-V // world)
    input_mb = per_gpu * H * W * (4 + 1 + 3) / 1e6
    return {"workload": workload, "views_per_gpu": per_gpu, "views_total": per_gpu * world if scaling == "weak" else V,
            "H": H, "W": W, "sfm_points_per_view": n_sfm, "aligner": aligner, "subsample": subsample,
            "patch_mask": True, "scaling": scaling,
            "l2": f"inputs {input_mb:.0f} MB per GPU and step > 126 MB L2, no flush needed"}


def cpu_reference_run(workload: str, n_views_sample: int, seed: int = 20261104):
    """The reference's CPU path (oracle port: same torch ops, same python loops, including the
    per-patch loop of num_sfm_points_mask) on a bounded sample of the workload's views."""
    import torch
    sys.path.insert(0, str(ROOT / "oracle"))
    import mdi_oracle as O

    V, H, W, n_sfm, aligner, subsample = WORKLOADS[workload]
    n = min(V, n_views_sample)
    torch.set_num_threads(os.cpu_count() or 1)
    sc = mod("synthetic").make_scene(n_views=n, H=H, W=W, n_sfm=n_sfm, seed=seed, device="cpu", low_conf_every=50)
    O.scene_points(sc, aligner=aligner, subsample=subsample, include_sfm_points=False, views=[0],
                   seed=1, slow_patch_loop=True)  # warm-up (thread pools, allocator)
    t0 = time.perf_counter()
    pts, rgbs, traces, kept = O.scene_points(sc, aligner=aligner, subsample=subsample, include_sfm_points=False,
                                             seed=1, slow_patch_loop=True)
    dt = time.perf_counter() - t0
    return n, dt, int(pts.shape[0]), torch.get_num_threads()


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    V, H, W, n_sfm, aligner, subsample = WORKLOADS[args.workload]
    times, kept_pts, nv, cores = [], 0, 0, 1
    for i in range(args.warmup + args.steps):
        nv, dt, kept_pts, cores = cpu_reference_run(args.workload, args.cpu_views, seed=20261104 + i)
        if i >= args.warmup:
            times.append(dt)
    t = sum(times) / max(len(times), 1)
    vps = nv / t
    sample = f"{nv} of {V} views of {args.workload} per step (views are independent; views/s is per view)"
    print(json.dumps({
        "impl": "reference", "metric": "init views/sec (align+subsample+unproject)", "value": vps, "unit": "views/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        # same config as our arm; the CPU run covers `run.sample_views_per_step` of the views per step
        "config": workload_config(args.workload, args.scaling, int(os.environ.get("WORLD_SIZE", "1"))),
        "run": {"sample_views_per_step": nv},
        "mpts_per_s": kept_pts / t / 1e6,
        "cpu_baseline": {"value": vps, "unit": "views/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": vps, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dmod = mod("distributed")
    # one process per GPU: run on (and allocate the pinned staging buffers from) the GPU's own NUMA node
    numa = dmod.bind_process_to_gpu_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = mod("_lib")
    mdi = mod("monocular_depth_init")
    emod = mod("engine")
    eng = emod.get_engine(dev)

    V_total, H, W, n_sfm, aligner, subsample = WORKLOADS[args.workload]
    if args.views:
        V_total = args.views
    strong = args.scaling == "strong"
    cfg = make_config(aligner, subsample)
    if strong:
        # BASELINE configs 3 / 4: ONE scene, its views block-partitioned over the ranks in dataset order.
        # A view is a pure function of (seed, view index), but the SfM table couples views (foreign
        # indices), so every rank generates the scene's small per-view data for all views and keeps the
        # depth / mask / image of its own block only.
        mine = dmod.view_partition(V_total, world, rank)
        sc_all = mod("synthetic").make_scene(n_views=V_total, H=H, W=W, n_sfm=n_sfm, seed=20261104, device=str(dev),
                                             low_conf_every=50, keep_views=mine)
        sc = sc_all
        V = len(mine)
    else:
        V = V_total
        sc = mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104 + 7919 * rank,
                                         device=str(dev), low_conf_every=50)
    depth, mask, image = sc.depth, sc.mask, sc.image   # resident in HBM before the timed region
    names, K, c2w = sc.names, sc.K, sc.c2w
    sfm = mdi.SfmTable(sc.points, sc.point_indices, names, dev)
    HW = H * W

    torch.manual_seed(1234)   # RANSAC device-sampler seeds are drawn from torch's CPU generator

    # N > 1: the exchange step (every rank ends up with the whole scene cloud) is fused into
    # the emit kernel over peer-mapped memory; NCCL all-gather only if symmetric memory is unavailable
    exchange = None
    gather = "none"
    if world > 1:
        try:
            exchange = dmod.CloudExchange(dev, capacity_rows=1 << 22, byte_rgb=os.environ.get("MDI_BYTE_RGB", "1" if world > 2 else "0") == "1")
            gather = ("emit kernel writes every rank's cloud through NVSwitch multicast (multimem.st.v4)" if exchange.use_multicast
                      else "emit kernel writes peer clouds over NVLink (16-byte stores, symmetric memory); counts exchanged on the device"
                           + ("; rgb travels as 3 bytes per point and is expanded by the receiver" if exchange.byte_rgb else ""))
        except Exception as e:  # noqa: BLE001 - any failure to map peer memory selects the NCCL path
            gather = f"nccl all_gather ({type(e).__name__})"

    def step(ex=exchange):
        res = mdi.points_from_views(cfg, depth, mask, image, K, c2w, names, sc.points,
                                    sc.point_indices, dev, L.IMAGE_U8, sfm=sfm, exchange=ex)
        if world > 1 and ex is None:
            res.xyz, res.rgb, _ = dmod.allgather_points(
                res.xyz if res.xyz is not None else torch.empty((0, 3), device=dev),
                res.rgb if res.rgb is not None else torch.empty((0, 3), device=dev))
        return res

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_loop(n_steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        r = None
        for _ in range(n_steps):
            r = step()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms), r

    if args.streams:
        emod.DEFAULT_STREAMS = args.streams
    n_warm = args.warmup if args.profile_mode else max(args.warmup, 3)
    for _ in range(n_warm):
        res = step()
    barrier()
    # ---- multi-GPU bit-identity, visible to the driver: the fused exchange cloud against the NCCL
    # all-gather of the per-rank clouds (same views, same RANSAC seeds), once, outside the timed region
    exchange_parity = None
    if world > 1 and exchange is not None and not args.profile_mode:
        torch.manual_seed(99)
        fused = step()
        fx, fr = fused.xyz.clone(), fused.rgb.clone()
        torch.manual_seed(99)
        plain = step(ex=None)
        ok = bool(fx.shape == plain.xyz.shape and torch.equal(fx, plain.xyz) and torch.equal(fr, plain.rgb))
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        exchange_parity = bool(int(flag))
        del fx, fr, fused, plain
        torch.manual_seed(1234)
        for _ in range(2):
            step()
        barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    lib = L.load()
    # timed loop 1: the same steps with CUDA events around every kernel launch of the library (per-kernel
    # times for the roofline objects); it also brings the clocks up for the headline loop
    # (one view range on one stream here, so that every kernel is timed alone)
    streams = emod.DEFAULT_STREAMS
    emod.DEFAULT_STREAMS = 1
    overlap = emod.OVERLAP_STATS
    emod.OVERLAP_STATS = False
    lib.mdi_profile_enable(1)
    L.profile_collect()
    ms_prof, res = timed_loop(args.steps)
    ktimes = L.profile_collect()
    # timed loop 2: the headline -- exactly K steps, no per-kernel events inside the timed region, the
    # batch queued as `streams` contiguous view ranges on separate CUDA streams (Engine.run)
    lib.mdi_profile_enable(0)
    emod.DEFAULT_STREAMS = streams
    emod.OVERLAP_STATS = overlap
    for _ in range(2):
        step()
    ms_total, res = timed_loop(args.steps)
    launches = sum(v["calls"] for k, v in ktimes.items() if k != "keep_zero_fill") // 1
    n_kept_total = int(res.xyz.shape[0]) if res.xyz is not None else 0
    n_kept = int(res.total)
    n_ok = int((res.status == 0).sum()) if res.status is not None else 0

    if args.profile_mode:
        if rank == 0:
            clocks.stop()
            print(json.dumps({"profile_mode": True, "streams": streams, "raw_stats_overlap": bool(overlap), "ms_per_step": ms_total / args.steps,
                              "ms_per_step_profiled": ms_prof / args.steps, "kernels": ktimes,
                              "kept": n_kept, "views": V}))
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- cold call: what a training run sees (the init path runs ONCE, runner.py:77-78): a fresh
    # Engine -- no cached descriptor table, no output-size hint, no workspaces -- kernels warm
    cold = []
    idx = dev.index
    for _ in range(3):
        emod._engines.pop(idx, None)
        barrier()
        t0 = time.perf_counter()
        step()
        torch.cuda.synchronize()
        cold.append(1e3 * (time.perf_counter() - t0))
    emod._engines[idx] = eng
    cold_ms = torch.tensor([sorted(cold)[1]], device=dev)
    if world > 1:
        dist.all_reduce(cold_ms, op=dist.ReduceOp.MAX)
    for _ in range(2):
        step()
    # ---- end to end: host buffers in, rgb out, through the public API --------------------------
    depth_h = depth.cpu().pin_memory(); mask_h = mask.cpu().pin_memory(); image_h = image.cpu().pin_memory()
    h2d = d2h = 0
    for _ in range(3):  # warm-up with the timed loop's own binding pattern: the previous result is still alive
        # while the next call allocates, so the caching allocator needs its second set of blocks before timing
        r2, rgb_h, h2d, d2h = mdi.points_from_host_views(cfg, depth_h, mask_h, image_h, K, c2w, names,
                                                         sc.points, sc.point_indices, dev, L.IMAGE_U8,
                                                         exchange=exchange)
        if world > 1 and exchange is None:
            dmod.allgather_points(r2.xyz, r2.rgb)
    barrier()
    t0 = time.perf_counter()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    e2e_steps = max(1, min(args.steps, 10))
    e2e_step_ms = []
    for _ in range(e2e_steps):
        t_step = time.perf_counter()
        r2, rgb_h, h2d, d2h = mdi.points_from_host_views(cfg, depth_h, mask_h, image_h, K, c2w, names,
                                                         sc.points, sc.point_indices, dev, L.IMAGE_U8,
                                                         exchange=exchange)
        if world > 1 and exchange is None:
            dmod.allgather_points(r2.xyz, r2.rgb)
        e2e_step_ms.append(round(1e3 * (time.perf_counter() - t_step), 2))
    f1.record()
    barrier()
    e2e_ms = torch.tensor([max(f0.elapsed_time(f1), 1e3 * (time.perf_counter() - t0))], device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    clk = clocks.stop() if rank == 0 else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    views_step = V_total if strong else world * V      # views all ranks process per step
    ms_step = ms_total / args.steps
    views_per_s = views_step / (ms_step * 1e-3)
    e2e_views_per_s = views_step / (float(e2e_ms) / e2e_steps * 1e-3)
    # ---- roofline of the dominant launch group ---------------------------------------------------
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    if peaks_path.exists():
        peak, peak_src = float(json.loads(peaks_path.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    n_slots = sum(len(sc.point_indices[n]) for n in names)
    adaptive = subsample == "adaptive"
    # with the exchange fused in, the emit kernel is bound by NVLink, not HBM: it gets its own roofline below
    hbm_calls = {k: v for k, v in ktimes.items() if algorithmic_bytes(k, HW, V, n_kept, n_slots, adaptive) > 0
                 and not (world > 1 and k == "emit_points_kernel")}
    dom = max(hbm_calls, key=lambda k: hbm_calls[k]["ms"]) if hbm_calls else None
    traffic = None   # dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu capture
    tpath = ROOT / "profiles" / "ncu_traffic.json"
    if tpath.exists() and dom:
        traffic = json.loads(tpath.read_text()).get(args.workload, {}).get(dom)
    roof = None
    breakdown = {}
    for k, v in ktimes.items():
        per = v["ms"] / v["calls"]
        b = algorithmic_bytes(k, HW, V, n_kept, n_slots, adaptive)
        breakdown[k] = {"ms_per_call": round(per, 4), "share": round(v["ms"] / ms_prof, 4),
                        "algorithmic_GBps": round(b / (per * 1e-3) / 1e9, 1) if b else None}
    if dom:
        per = ktimes[dom]["ms"] / ktimes[dom]["calls"]
        ach = algorithmic_bytes(dom, HW, V, n_kept, n_slots, adaptive) / (per * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "peak_source": peak_src,
                "note": "achieved = algorithmic bytes of one launch / its mean CUDA-event duration inside the (profiled) timed loop"}
    # N > 1: every GPU receives the other ranks' rows (24 B each) through the emit kernels' peer stores
    roof_exchange = None
    if world > 1 and "emit_points_kernel" in ktimes and exchange is not None:
        per = ktimes["emit_points_kernel"]["ms"] / ktimes["emit_points_kernel"]["calls"]
        ingress = (15.0 if exchange.byte_rgb else 24.0) * (n_kept_total - n_kept)
        ach = ingress / (per * 1e-3) / 1e9
        roof_exchange = {"bound": "nvlink", "kernel": "emit_points_kernel", "ms": round(per, 4),
                         "ingress_bytes_per_gpu": int(ingress), "achieved": ach, "peak": 770.0, "unit": "GB/s",
                         "frac": ach / 770.0,
                         "peak_source": "measured peer copy per direction on this pool (B200_PROFILING.md; nominal 900)"}
    # the a7 statistics group as a whole: ONE algorithmic read of depth + mask over the sum of its kernels
    group = [k for k in ktimes if k.startswith("stats_")]
    roof_group = None
    if group and adaptive:
        g_ms = sum(ktimes[k]["ms"] / ktimes[k]["calls"] for k in group)
        g_ach = V * HW * 5.0 / (g_ms * 1e-3) / 1e9
        roof_group = {"bound": "hbm", "kernels": sorted(group), "ms": round(g_ms, 4), "achieved": g_ach, "peak": peak,
                      "unit": "GB/s", "frac": g_ach / peak}
    # whole-step algorithmic bytes: P*HW*5 + HW + 36*Nk + 16*M per view
    P = 2 if adaptive else 1
    step_bytes = V * (P * HW * 5 + HW) + 36 * n_kept + 16 * n_slots
    nv, dt, kept_cpu, cores = cpu_reference_run(args.workload, args.cpu_views)
    out = {
        "metric": "init views/sec (align+subsample+unproject)", "value": views_per_s, "unit": "views/s",
        "n_gpus": world, "steps": args.steps, "warmup": n_warm, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, args.scaling, world),
        "run": {"views_ok_rank0": n_ok, "kept_points_rank0": n_kept, "cloud_rows": n_kept_total, "exchange": gather,
                "numa": numa, "streams": streams, "raw_stats_overlap": bool(overlap),
                "streams_note": "what the headline loop used: `streams` view ranges on separate CUDA streams, and "
                                "(raw_stats_overlap) the a7 statistics pass on the raw depth beside the alignment kernels; "
                                "the profiled loop (per-kernel times, roofline) runs every kernel alone on one stream"},
        "ms_per_step_profiled": ms_prof / args.steps,
        "cold_call_ms": float(cold_ms),
        "cold_call_note": "one call on a fresh Engine (no cached descriptor table, no output-size hint, no workspaces), "
                          "kernels warm, host wall clock to completion; median of 3, max over ranks",
        "exchange_parity": exchange_parity,
        "mpts_per_s": (n_kept_total if world > 1 else n_kept) / (ms_step * 1e-3) / 1e6,
        "step_algorithmic_GBps": step_bytes / (ms_step * 1e-3) / 1e9,
        "step_frac_of_peak": step_bytes / (ms_step * 1e-3) / 1e9 / peak,
        "roofline": roof, "roofline_group": roof_group, "roofline_exchange": roof_exchange, "kernels": breakdown,
        "cpu_baseline": {"value": nv / dt, "unit": "views/s", "cores": cores, "kind": "port",
                         "sample": f"{nv} of {V_total} views of the same workload, oracle port of the reference loop on the host"},
        "e2e": {"value": e2e_views_per_s, "unit": "views/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "step_ms": e2e_step_ms},
        "gpu_launches": launches, "clocks": clk,
    }
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def run_postprocess(args):
    """C5 post-processing leg (a12-a14): the hot path runs once per rank and the exchange leaves the
    whole scene cloud on every GPU; then K timed steps of LOF outlier removal (kNN k = 40 + LOF, queries
    split over the ranks) or of the native merge subsampling (extents split over the ranks, the k-d merge
    on every rank).  value = points of the scene cloud per second."""
    import numpy as np
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dmod = mod("distributed")
    dmod.bind_process_to_gpu_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L, mdi, knn = mod("_lib"), mod("monocular_depth_init"), mod("knn")
    sub, pcfg = mod("point_cloud_postprocess.subsampling"), mod("point_cloud_postprocess.config")
    V_total, H, W, n_sfm, aligner, subsample = WORKLOADS[args.workload]
    V = args.views or max(1, V_total // 8)          # C5: 2000 views on 8 GPUs = 250 per GPU
    cfg = make_config(aligner, subsample)
    sc = mod("synthetic").make_scene(n_views=V, H=H, W=W, n_sfm=n_sfm, seed=20261104 + 7919 * rank, device=str(dev),
                                     low_conf_every=50)
    torch.manual_seed(1234)
    exchange = dmod.CloudExchange(dev, capacity_rows=1 << 22) if world > 1 else None
    res = mdi.points_from_views(cfg, sc.depth, sc.mask, sc.image, sc.K, sc.c2w, sc.names, sc.points, sc.point_indices,
                                dev, L.IMAGE_U8, exchange=exchange)
    pts, rgb = res.xyz.clone(), res.rgb.clone()
    ok = (res.status == 0)
    Kd, Pd = sc.K.to(dev)[ok], res.P[ok]
    sizes = torch.tensor([[W, H]], dtype=torch.int32, device=dev).repeat(int(ok.sum()), 1)
    if world > 1:   # cameras of all ranks (a13 needs every camera): padded all-gather, rank order = view order
        def gather_cams(t):
            n = torch.tensor([t.shape[0]], device=dev)
            ns = [torch.zeros_like(n) for _ in range(world)]
            dist.all_gather(ns, n)
            m = int(max(x.item() for x in ns))
            pad = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
            pad[: t.shape[0]] = t
            outs = [torch.empty_like(pad) for _ in range(world)]
            dist.all_gather(outs, pad)
            return torch.cat([o[: int(k.item())] for o, k in zip(outs, ns)])
        Kd, Pd, sizes = gather_cams(Kd), gather_cams(Pd), gather_cams(sizes)
    del sc, res, exchange
    torch.cuda.empty_cache()
    n = int(pts.shape[0])
    group = dist.group.WORLD if world > 1 else None
    params = pcfg.PointCloudSubsamplingParams()

    def step():
        if args.postprocess == "lof":
            out, _ = (knn.local_outlier_factor_sharded(pts, 40, group) if group is not None
                      else knn.local_outlier_factor(pts, 40))
            keep = ~out
            return pts[keep], rgb[keep]
        p, c, _ = sub.subsample_pointcloud_device(pts, rgb, Kd, Pd, sizes, params, group=group)
        return p, c

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n_warm = max(args.warmup, 1)
    for _ in range(n_warm):
        out = step()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    lib = L.load()
    lib.mdi_profile_enable(1)
    L.profile_collect()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        out = step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ktimes = L.profile_collect()
    lib.mdi_profile_enable(0)
    # every rank must end with the same cloud (bit-identity itself: tests/postprocess_check.py)
    same = None
    if world > 1:
        flag = torch.tensor([int(out[0].shape[0])], device=dev)
        lo, hi = flag.clone(), flag.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(int(lo) == int(hi))
    clk = clocks.stop() if rank == 0 else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ms_step = float(ms) / args.steps
    # ---- CPU beside it: sklearn (LOF) / the reference's own compiled module (merge) on a bounded subset ----
    pts_h, rgb_h = pts.cpu().numpy(), rgb.cpu().numpy()
    if args.postprocess == "lof":
        from sklearn.neighbors import LocalOutlierFactor
        m = min(n, args.cpu_points)
        t0 = time.perf_counter()
        LocalOutlierFactor(n_neighbors=40, n_jobs=-1).fit_predict(pts_h[:m])
        dt = time.perf_counter() - t0
        cpu = {"value": m / dt / 1e6, "unit": "Mpts/s", "cores": os.cpu_count(), "kind": "reference",
               "sample": f"sklearn LocalOutlierFactor(n_neighbors=40, n_jobs=-1) -- the reference's call, postprocess.py:20 -- "
                         f"on the first {m} of {n} points"}
    else:
        sys.path.insert(0, str(ROOT / "oracle"))
        m = min(n, args.cpu_points)
        cams = min(int(Kd.shape[0]), 200)
        Kh, Ph, sh = Kd[:cams].cpu().numpy(), Pd[:cams].cpu().numpy(), sizes[:cams].cpu().numpy()
        kind = "port"
        try:
            import build_ref_native
            ref = build_ref_native.load() if build_ref_native.built() else None
        except Exception:  # noqa: BLE001
            ref = None
        t0 = time.perf_counter()
        if ref is not None:
            kind = "reference"
            from types import SimpleNamespace
            cwd = os.getcwd(); os.chdir("/tmp")
            devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1); os.dup2(devnull, 1)
            try:
                ref.subsample_pointcloud(pts_h[:m], rgb_h[:m], list(Kh), list(Ph), sh,
                                         SimpleNamespace(max_bbox_aspect_ratio=1.1, min_extent_multiplier=1.0))
            finally:
                os.dup2(saved, 1); os.close(devnull); os.close(saved); os.chdir(cwd)
        else:
            import native_oracle as NO
            NO.subsample_pointcloud(pts_h[:m], rgb_h[:m], Kh, Ph, sh)
        dt = time.perf_counter() - t0
        cpu = {"value": m / dt / 1e6, "unit": "Mpts/s", "cores": 1, "kind": kind,
               "sample": f"{'the reference module compiled into oracle/_ref' if kind == 'reference' else 'C restatement'} "
                         f"(single-threaded like the original) on the first {m} of {n} points, {cams} of {int(Kd.shape[0])} cameras"}
    print(json.dumps({
        "metric": f"postprocess Mpts/sec ({args.postprocess}: " + ("kNN k=40 + LOF + filter" if args.postprocess == "lof" else "min extents + k-d merge") + ")",
        "value": n / (ms_step * 1e-3) / 1e6, "unit": "Mpts/s", "n_gpus": world, "steps": args.steps, "warmup": n_warm,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": args.workload, "postprocess": args.postprocess, "views_per_gpu": V, "H": H, "W": W,
                   "cloud_points": n, "cameras": int(Kd.shape[0]),
                   "split": "kNN / LOF queries and extents split over the ranks, 4 + 4 + 8 B per point exchanged; "
                            "k-d merge on every rank (replicas only)"},
        "kept_points": int(out[0].shape[0]), "same_result_on_every_rank": same,
        "kernels": {k: {"ms_per_call": round(v["ms"] / v["calls"], 4), "calls_per_step": v["calls"] / args.steps}
                    for k, v in ktimes.items()},
        "cpu_baseline": cpu, "clocks": clk,
    }))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS))
    ap.add_argument("--cpu-views", type=int, default=12, help="views in the bounded CPU sample")
    ap.add_argument("--views", type=int, default=0, help="override views per GPU (profiling runs only)")
    ap.add_argument("--postprocess", default=None, choices=["lof", "merge"],
                    help="C5 post-processing leg instead of the hot path (default workload: c5_large_...)")
    ap.add_argument("--cpu-points", type=int, default=1000000, help="points in the bounded CPU sample of --postprocess")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank owns the workload's views; strong: the workload's views are sharded over the ranks")
    ap.add_argument("--streams", type=int, default=0,
                    help="view ranges of a step queued on separate CUDA streams (0: the engine default, MDI_STREAMS)")
    ap.add_argument("--profile-mode", action="store_true",
                    help="short run for ncu: warm-up as given, no e2e / CPU legs (not a bench value)")
    args = ap.parse_args()
    if args.postprocess and args.workload == DEFAULT_WORKLOAD:
        args.workload = "c5_large_2000x1600x1200_ransac_adaptive"
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.postprocess:
        run_postprocess(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
