#!/usr/bin/env python
"""bench.py -- rays/sec of DM-NeRF's per-ray hot path on B200 (BASELINE.json metric).

Workload at every N: DM-SR full-image render (BASELINE.json configs[1]): 640x480 = 307,200 rays,
64 coarse + 128 fine samples per ray, K=32 instances, RGB + depth + instance map (+ uint8 labels),
synthetic pose, random-init weights.  One "step" = one full frame per rank (weak scaling: each rank
renders its own view of the sweep; rank 0 gathers the label maps).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode bf16|fp32] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (rays already in HBM),
`e2e` = same frame through the public API with pinned HOST ray buffers in and host results out,
`roofline` = the dominant kernel (fused PE+MLP, fine pass) against the measured bf16 tensor peak,
`cpu_baseline` = the CPU oracle port (torch CPU fp32, all host threads) on a bounded ray sample.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, K_INS, SC, NIMP = 480, 640, 32, 64, 128
NEAR, FAR = 4.0, 15.0
N_RAYS = H * W
CPU_SAMPLE_RAYS = 24576                          # ~10-20 s of host work for the cpu_baseline leg
MACS_PER_POINT = 691840 + 128 * K_INS            # algorithmic, padding not counted
FLOP_PER_RAY = 2 * MACS_PER_POINT * (SC + SC + NIMP)


def synth_camera(torch, view: int):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    focal = 0.5 * W / math.tan(0.5 * 0.6911)
    Kmat = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]], dtype=torch.float32)
    th, ph, rad = (37.0 + 45.0 * view) / 180 * math.pi, -65.0 / 180 * math.pi, 7.0
    t = torch.eye(4); t[2, 3] = rad
    rp = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0], [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1]])
    rt = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0], [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1]])
    flip = torch.tensor([[-1., 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]])
    return Kmat, (flip @ rt.float() @ rp.float() @ t)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self._stop_evt.is_set():
            try:
                r = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5)
                if r.returncode == 0 and r.stdout.strip():
                    self.rows.append([c.strip() for c in r.stdout.strip().split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def summary(self):
        self._stop_evt.set()
        self.join(timeout=3)
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        reasons = []
        for i, name in enumerate(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")):
            if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        mx = max((int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons, "samples": len(sm)}


def cpu_oracle_rays_per_s(torch, n_rays: int, repeats: int = 1):
    """The oracle port (torch CPU fp32, reference op order) on a bounded sample of the same workload:
    n_rays of the frame, tester.py-style chunk.  Returns (rays/s, cores, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dmnerf_oracle as orc
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pc, pf = orc.init_params(K_INS, 1), orc.init_params(K_INS, 2)
    Kmat, c2w = synth_camera(torch, 0)
    ro, rd = orc.get_rays_k(H, W, Kmat, c2w)
    sel = torch.linspace(0, N_RAYS - 1, n_rays).long()
    rays = torch.stack([ro.reshape(-1, 3)[sel], rd.reshape(-1, 3)[sel]], 0)
    zc = orc.z_val_sample(n_rays, NEAR, FAR, SC)
    best = None
    with torch.no_grad():
        for _ in range(repeats):
            t0 = time.perf_counter()
            for a in range(0, n_rays, 4096):                       # tester.py:62 chunking (N_test = 4096)
                orc.dm_nerf(rays[:, a:a + 4096], pc, pf, zc[a:a + 4096], perturb=0.0, n_importance=NIMP)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
    return n_rays / best, cores, best


def cpu_oracle_train_rays_per_s(torch, n_rays: int):
    """One training step of the oracle port on the host cores (forward, MSE + penalizer losses, backward)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dmnerf_oracle as orc
    torch.set_num_threads(os.cpu_count() or 1)
    pc = {k: v.requires_grad_(True) for k, v in orc.init_params(K_INS, 1).items()}
    pf = {k: v.requires_grad_(True) for k, v in orc.init_params(K_INS, 2).items()}
    Kmat, c2w = synth_camera(torch, 0)
    ro, rd = orc.get_rays_k(H, W, Kmat, c2w)
    sel = torch.linspace(0, N_RAYS - 1, n_rays).long()
    rays = torch.stack([ro.reshape(-1, 3)[sel], rd.reshape(-1, 3)[sel]], 0)
    zc = orc.z_val_sample(n_rays, NEAR, FAR, SC)
    tgt = torch.rand(n_rays, 3)
    t0 = time.perf_counter()
    o = orc.dm_nerf(rays, pc, pf, zc, perturb=1.0, n_importance=NIMP, t_rand=torch.rand(n_rays, SC),
                    u=torch.rand(n_rays, NIMP))
    loss = ((o["rgb_fine"] - tgt) ** 2).mean() + ((o["rgb_coarse"] - tgt) ** 2).mean()
    loss = loss + orc.ins_penalizer(o["raw_coarse"], o["z_vals_coarse"], o["depth_coarse"], rays[1], 0.05, 0.05)
    loss = loss + orc.ins_penalizer(o["raw_fine"], o["z_vals_fine"], o["depth_fine"], rays[1], 0.05, 0.05)
    loss.backward()
    dt = time.perf_counter() - t0
    return n_rays / dt, dt


def run_reference(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = 8192
    times = []
    for i in range(args.warmup + args.steps):
        rps, cores, dt = cpu_oracle_rays_per_s(torch, n)
        if i >= args.warmup:
            times.append(dt)
    ms = 1000 * sum(times) / len(times)
    val = n / (ms / 1000)
    sample = f"{n} rays of the 640x480 frame per step (oracle port, torch CPU fp32, {cores} threads)"
    emit({
        "impl": "reference", "metric": "rays/sec render (64+128 samples/ray)", "value": val, "unit": "rays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "DM-SR full-image render 640x480, 64+128 samples/ray, K=32 (bounded sample)"},
        "cpu_baseline": {"value": val, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def train_leg(args, torch, dist, dev, rank, world):
    """Training throughput (BASELINE.json configs[0] and configs[2]): full steps of dm_nerf_b200.train.TrainStep --
    pixel draw, dm_nerf fwd, MSE + Hungarian instance loss + emptiness penalizer, backward, one all-reduce of the
    flat gradient, Adam -- data-parallel over ray batches (each rank its own view and rays; weak scaling)."""
    import numpy as np
    import dmnerf_oracle as orc
    from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
    from dm_nerf_b200.train import TrainStep
    pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
    out = {}
    for tag, n, k, near, far in (("C1_study_1024rays_K32", 1024, 32, 4.0, 15.0), ("C3_replica_4096rays_K93", 4096, 93, 0.0, 6.5)):
        torch.manual_seed(3 + rank); np.random.seed(rank)
        mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
        mc.load_state_dict(orc.init_params(k, 1)); mf.load_state_dict(orc.init_params(k, 2))
        mc, mf = mc.to(dev), mf.to(dev)
        Kmat, c2w = synth_camera(torch, rank)
        pose = c2w[:3, :4].to(dev)
        img = torch.rand(H, W, 3, device=dev)
        lab = ((torch.arange(H, device=dev)[:, None] // 60) * 8 + (torch.arange(W, device=dev)[None] // 80)) % k
        targs = SimpleNamespace(N_train=n, near=near, far=far, N_samples=SC, N_importance=NIMP, perturb=1.0, ins_num=k,
                                tolerance=0.05, deta_w=0.05, lrate=5e-4, lrate_decay=500)
        # the whole step is one CUDA-graph replay (device-side Hungarian, no host round trip); inputs are copied
        # into the graph's static buffers every step, as a real loop does with the image it drew
        ts = TrainStep(pe, ve, mc, mf, Kmat, targs, penalize=True, graph=True)
        for _ in range(max(3, args.warmup)):
            ts.step(img, pose, lab)
        steps = max(10, args.steps)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            losses = ts.step(img, pose, lab)
        b.record()
        torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(b) / steps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms)
        macs = 691840 + 128 * k
        flop_per_ray = 2.0 * (3 * macs - 101248) * (SC + SC + NIMP)            # fwd + dgrad + wgrad
        out[tag] = {"value": n * world / (ms / 1000), "unit": "rays/s", "ms_per_step": ms, "rays_per_step_per_gpu": n,
                    "steps": steps, "loss": float(losses["total"]),
                    "algorithmic_tflops": flop_per_ray * n * world / (ms / 1000) / 1e12}
    return {"metric": "rays/sec train (64+128 samples/ray, fwd+bwd+Adam, penalizer on)", "scaling": "weak",
            "step": "dm_nerf_b200.train.TrainStep(graph=True): one CUDA-graph replay per step",
            "parallelism": f"data-parallel ray batches x{world}, one all-reduce of the flat gradient per step",
            "configs": out}


def run_gpu(args):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dmnerf_oracle as orc
    from dm_nerf_b200 import ops, _lib
    from dm_nerf_b200.frame import render_pixels
    from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
    from dm_nerf_b200.parallel import gather_rows

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    mode = args.mode
    if mode == "bf16" and _lib.lib().dmnerf_packed_bytes(K_INS, _lib.MODE_BF16) <= 0:
        raise RuntimeError("bf16 tensor-core mode is not built")
    ops.set_mode(mode)
    mc, mf = DM_NeRF(8, 256, 63, 27, [4], K_INS), DM_NeRF(8, 256, 63, 27, [4], K_INS)
    mc.load_state_dict(orc.init_params(K_INS, 1)); mf.load_state_dict(orc.init_params(K_INS, 2))
    mc, mf = mc.to(dev), mf.to(dev)
    pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
    Kmat, c2w = synth_camera(torch, rank)                      # each rank renders its own view of the sweep
    hwk = (H, W, Kmat)
    ro, rd = ops.rays_k(H, W, Kmat, c2w, 0, N_RAYS, dev)
    rays_dev = torch.stack([ro, rd], 0)
    rays_host = rays_dev.cpu().pin_memory()
    out = {"rgb": torch.empty((N_RAYS, 3), device=dev), "depth": torch.empty((N_RAYS,), device=dev),
           "ins": torch.empty((N_RAYS, K_INS), device=dev), "labels": torch.empty((N_RAYS,), dtype=torch.uint8, device=dev)}
    host_out = {k: torch.empty(v.shape, dtype=v.dtype).pin_memory() for k, v in out.items() if k != "ins"}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def frame(rays):
        render_pixels(pe, ve, mc, mf, hwk, c2w, NEAR, FAR, SC, NIMP, rays=rays, chunk=args.chunk, out=out)
        if world > 1:
            gather_rows(out["labels"], N_RAYS * world)          # the final gather (rank 0 collects label maps)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """max-over-ranks device time (ms) of `steps` calls, CUDA events on the launching stream."""
        evs = []
        sync()
        for _ in range(steps):
            flush.zero_()                                        # L2 flush between timed iterations
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            evs.append((a, b))
        sync()
        ms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps

    for _ in range(args.warmup):
        frame(rays_dev)
    sync()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = ops.launches
    with ops.KernelTimer() as kt:
        ms_step = timed(lambda: frame(rays_dev), args.steps)
    launches = (ops.launches - l0)
    totals = kt.totals()

    def e2e_step():
        r = rays_host.to(dev, non_blocking=True)
        frame(r)
        for k in host_out:
            host_out[k].copy_(out[k], non_blocking=True)
    e2e_step(); sync()
    ms_e2e = timed(e2e_step, args.steps)
    clocks = sampler.summary()
    train = None
    if not args.no_train:
        train = train_leg(args, torch, dist, dev, rank, world)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    mode_id = _lib.MODE_BF16 if mode == "bf16" else _lib.MODE_FP32
    name = f"dmnerf_mlp_fwd[S={SC + NIMP},mode={mode_id}]"
    n_launch, ms_kernel = totals.get(name, (0, 0.0))
    rays_per_launch = min(args.chunk, N_RAYS)
    roofline = None
    if n_launch:
        traffic = ncu_traffic_bytes() if (mode == "bf16" and rays_per_launch == 32768 and K_INS == 32) else None
        # achieved = algorithmic FLOPs per launch / mean launch duration (fine pass: 192 points per ray)
        flops_total = 2.0 * MACS_PER_POINT * (SC + NIMP) * N_RAYS * args.steps
        achieved = flops_total / (ms_kernel / 1000) / 1e12
        peak = peaks.get("bf16_tflops_sustained", 1400.0)
        roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                    "traffic": traffic, "traffic_source": "profiles/r01_render_kernels_metrics.csv (ncu --set full, "
                    "dram__bytes_read.sum + dram__bytes_write.sum of the fine-pass launch)" if traffic else None,
                    "kernel": name, "launches": n_launch, "ms_per_launch": ms_kernel / n_launch,
                    "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback",
                    "kernel_share_of_step": ms_kernel / (ms_step * args.steps)}
    stage_ms = {k: round(v[1] / args.steps, 3) for k, v in totals.items()}
    # the HBM-bound ray kernels of the same frame against the measured copy bandwidth (algorithmic bytes per
    # ray, K=32: compositing fine 29,340 / coarse 10,140; sampler 1,280 with in-kernel linspace uniforms)
    hbm_peak = peaks.get("hbm_gbs", 6400.0)
    secondary = []
    for kname, bytes_per_ray in ((f"dmnerf_composite_fwd[S={SC + NIMP}]", 29340), (f"dmnerf_composite_fwd[S={SC}]", 10140),
                                 ("dmnerf_fine_z", 1280)):
        cnt, ms_k = totals.get(kname, (0, 0.0))
        if cnt:
            gbps = bytes_per_ray * N_RAYS * args.steps / (ms_k / 1000) / 1e9
            secondary.append({"kernel": kname, "bound": "hbm" if "composite" in kname else "latency (CDF + searches per ray; HBM figure for scale)", "achieved": gbps, "peak": hbm_peak, "unit": "GB/s",
                              "frac": gbps / hbm_peak, "ms_per_launch": ms_k / cnt})

    if rank == 0:
        cpu = None
        try:
            rps, cores, secs = cpu_oracle_rays_per_s(torch, CPU_SAMPLE_RAYS)
            cpu = {"value": rps, "unit": "rays/s", "cores": cores, "kind": "port",
                   "sample": f"{CPU_SAMPLE_RAYS} of the frame's 307200 rays in tester.py-style 4096-ray chunks, one pass "
                             f"({secs:.1f} s), oracle port torch CPU fp32"}
            if train is not None:
                t_rps, t_secs = cpu_oracle_train_rays_per_s(torch, 1024)
                train["cpu_baseline"] = {"value": t_rps, "unit": "rays/s", "cores": cores, "kind": "port",
                                         "sample": f"one C1 step (1024 rays, K=32: dm_nerf + MSE + penalizer + backward, "
                                                   f"{t_secs:.1f} s), oracle port torch CPU fp32"}
        except Exception as e:  # pragma: no cover
            cpu = {"value": None, "error": str(e)}
        total_rays = N_RAYS * world
        h2d = rays_host.numel() * 4
        d2h = sum(v.numel() * v.element_size() for v in host_out.values())
        emit({
            "metric": "rays/sec render (64+128 samples/ray)", "value": total_rays / (ms_step / 1000), "unit": "rays/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if mode == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": "DM-SR full-image render 640x480 = 307200 rays x (64+128) samples, K=32, "
                                   "rgb+depth+instance map+labels, one frame per rank per step",
                       "mode": mode, "chunk_rays": args.chunk, "l2": "flushed between timed iterations (256 MiB write)",
                       "parallelism": f"rays/views sharded x{world}, final gather of label maps"},
            "e2e": {"value": total_rays / (ms_e2e / 1000), "unit": "rays/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e},
            "gpu_launches": launches, "roofline": roofline, "roofline_secondary": secondary, "cpu_baseline": cpu, "clocks": clocks,
            "stage_ms_per_step": stage_ms, "train": train,
        })
    if world > 1:
        dist.destroy_process_group()


def ncu_traffic_bytes():
    """DRAM bytes per launch of the dominant kernel (fine-pass MLP forward, 32768 rays x 192 points, K=32) from the
    committed ncu --set full capture of this same workload; None when the summary is absent."""
    try:
        rows = {}
        with open(os.path.join(ROOT, "profiles", "r01_render_kernels_metrics.csv")) as f:
            for line in f:
                if line.startswith("#"):
                    continue
                parts = line.rstrip("\n").split(",")
                rows[parts[0]] = parts[1:]
        names, dur = rows["Kernel Name"][1:], [float(x) for x in rows["gpu__time_duration.sum"][1:]]
        idx = max((i for i, nm in enumerate(names) if "mlp_fwd_bf16" in nm), key=lambda i: dur[i])
        scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
        total = 0.0
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            total += float(rows[key][1:][idx]) * scale[rows[key][0]]
        return total
    except Exception:
        return None


_JSON_FD = None


def quiet_stdout():
    """Send everything libraries write to fd 1 (NCCL prints its version banner there) to stderr; the one JSON line
    goes to the saved descriptor through emit()."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    if _JSON_FD is None:
        os.write(1, line)
    else:
        os.write(_JSON_FD, line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default=os.environ.get("DMNERF_BENCH_MODE", "auto"), choices=["auto", "bf16", "fp32"])
    ap.add_argument("--chunk", type=int, default=32768)
    ap.add_argument("--no-train", action="store_true", help="skip the training-throughput leg")
    args = ap.parse_args()
    if not (args.gpus > 1 and int(os.environ.get("WORLD_SIZE", "1")) == 1 and args.impl != "reference"):
        quiet_stdout()                                  # (the torchrun relaunch path leaves fd 1 to its children)
    if args.impl == "reference":
        return run_reference(args)
    if args.mode == "auto":
        from dm_nerf_b200 import _lib
        args.mode = "bf16" if _lib.lib().dmnerf_packed_bytes(K_INS, _lib.MODE_BF16) > 0 else "fp32"
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: relaunch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    run_gpu(args)


if __name__ == "__main__":
    main()
