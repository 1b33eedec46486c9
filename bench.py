#!/usr/bin/env python
"""bench.py — Mpaths/s of the B200 wavefront path tracer on BASELINE.json's workload, one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A step = one render of the workload image (every path of every pixel) = the `for y / for x / for spp` nest of
the reference's example drivers (examples/cornell-box.rs:155-181).  Default workload: BASELINE.json configs[1],
cornell-box 600x600, 1024 spp, depth 50 (368.6 M paths per step).

  value      Mpaths/s, whole job over all N GPUs, scene resident in HBM, framebuffer left in HBM (rt_render_device
             + for N>1 the NCCL all-gather and the de-interleave kernel), timed with CUDA events on the stream the
             kernels run on, max over ranks.  N>1 shards the SAME image by interleaved rows -> "scaling": "strong".
  e2e        same metric through the host-pointer C-ABI calls a reference user makes (what rust/src/gpu.rs binds): every step
             rt_scene_create (flat scene tables from host memory: device BVH build + H2D) + rt_render (host framebuffer out,
             D2H inside) at N = 1; at N > 1 the same through rt_render_device + NCCL gather + D2H on rank 0.
  roofline   the DOMINANT kernel of the step — whichever stage kernel the live CUDA events say takes longest (cornell-box:
             k_shade<2>, HBM bound; book-one: k_extend<0,1>, ALU-issue bound) — in the contract's HBM form (algorithmic queue
             bytes per ray / launch time); `roofline_fp32` / `roofline_l2` state extend's instruction-issue and L2 fractions.
  book_one_4k  the north-star scene (BASELINE.json configs[4], book-one 3840x2160, 1024 spp, depth 50) measured in the SAME run
             with the same method, so that the driver's BENCH / SCALE records time it at every N.
  multi_c_abi  at N > 1: the same image through rt_render_multi — ONE process, one host thread per device, ncclCommInitAll,
             gather to device 0 — driven by rank 0 after the per-rank measurement (the other ranks idle at a barrier).
  cpu_baseline  the f64 oracle in FAITHFUL mode (the reference's algorithm: unpruned recursive BVH, Mat4 per sprite,
             rejection sampling), one thread per host core, on a bounded sample of the same image (fewer spp).
--impl reference times that CPU implementation as its own arm (the Rust reference cannot be built in this image).
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"  # keep NCCL's version banner off stdout: rank 0 prints exactly one JSON line

WORKLOADS = {
    # name: (scene, width, height, spp, depth, BASELINE.json config it is)
    "cornell-box": ("cornell-box", 600, 600, 1024, 50, "configs[1] examples/cornell-box.rs 600x600 1024spp depth50"),
    "book-one": ("book-one", 400, 225, 100, 50, "configs[0] examples/book-one.rs 400x225 100spp depth50"),
    "book-two": ("book-two", 800, 800, 1024, 50, "configs[2] examples/main.rs 800x800 1024spp depth50, synthetic 1024x512 earth texture"),
    "stress-1m": ("stress:1000000", 3840, 2160, 256, 16, "configs[3] 1M random spheres 3840x2160 256spp depth16"),
    "book-one-4k": ("book-one", 3840, 2160, 1024, 50, "configs[4] book-one 3840x2160 1024spp depth50"),
}
# algorithmic HBM bytes per ray segment of k_extend: reads a (16 B) + b (16 B) of the ray queue, writes the hit (8 B)
EXTEND_BYTES_PER_RAY = 40
# SURVEY §8(d) per-unit FP32 work of the extend kernel (lane-instructions)
FP32_PER_BOX, FP32_PER_SPHERE, FP32_PER_RECT = 20, 30, 35


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(p.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self, wait_s=4.0):
        """Starts nvidia-smi (one row every 50 ms) and waits for its first row, so that even a timed region of a few tens of
        milliseconds (8 GPUs: 10 ms per step) has the sampler running."""
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        threading.Thread(target=self._read, daemon=True).start()
        t_end = time.perf_counter() + wait_s
        while not self.rows and time.perf_counter() < t_end:
            time.sleep(0.01)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([time.perf_counter()] + [c.strip() for c in line.split(",")])

    def stop(self, t0=None, t1=None, t2=None):
        """Clocks over the rows that arrived in [t0, t1] (the timed region); when the region was too short to catch one, over [t0, t2]
        (the region plus the untimed single-lane repeat of the same steps that follows it) — `window` says which."""
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        window = "all samples"
        rows = self.rows
        if t0 is not None:
            rows = [r for r in self.rows if t0 <= r[0] <= t1 + 0.05]
            window = "timed region"
            if not rows and t2 is not None:
                rows = [r for r in self.rows if t0 <= r[0] <= t2 + 0.05]
                window = "timed region + the untimed repeat of the same steps right after it"
        self.rows = [r[1:] for r in rows]
        out = self._summary()
        out["window"] = window
        return out

    def _summary(self):
        sm, reasons, mx = [], set(), None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for name, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": mx, "reasons": [], "samples": 0}
        busy = sorted(sm)[len(sm) // 4:]  # drop the idle tail of the first samples
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def measure_l2():
    """L2 read-stream bandwidth measured on this box by ray-tracer_b200/tools/l2bw (no L2 figure in MEASURED_PEAKS.json)."""
    tool = os.path.join(ROOT, "ray-tracer_b200", "tools", "l2bw")
    if not os.path.exists(tool):
        return None
    try:
        out = subprocess.run([tool], capture_output=True, text=True, timeout=120).stdout.strip().splitlines()[-1]
        return json.loads(out)
    except Exception:  # noqa: BLE001 - a missing figure is reported as null, never guessed
        return None


def ncu_traffic(workload, kernel):
    """DRAM bytes per ray of a kernel from the committed `ncu --set full` capture of the same workload (profiles/traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum of one launch / its rays).  Not measured in this run — hardware counters
    need the profiler — so the line says so (`traffic_measured_in_run: false`) next to the capture it comes from."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None, None
    for t in json.load(open(path)).get(workload, []):
        if t.get("kernel") == kernel:
            return float(t["dram_bytes_per_ray"]), t.get("source")
    return None, None


def run_reference(args, wl):
    """The reference's own CPU implementation of the path, restated (oracle FAITHFUL mode), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as orc  # the one place besides cpu_baseline where bench.py executes oracle/

    scene_name, w, h, spp, depth, cfg = wl
    cores = os.cpu_count() or 1
    s = orc.Scene(scene_name, 1)
    # bounded sample: the same image at a few spp (throughput in paths/s does not depend on spp)
    probe_spp = 1
    t0 = time.time()
    _, _, st = s.render(w, h, probe_spp, depth, seed=1, mode=orc.FAITHFUL, threads=cores)
    rate = st.paths / max(time.time() - t0, 1e-6)
    budget_s = 150.0 / max(args.steps + args.warmup, 1)
    step_spp = max(1, min(spp, int(rate * min(budget_s, 20.0) / (w * h))))
    for i in range(args.warmup):
        s.render(w, h, step_spp, depth, seed=100 + i, mode=orc.FAITHFUL, threads=cores)
    paths = rays = 0
    t0 = time.time()
    for i in range(args.steps):
        _, _, st = s.render(w, h, step_spp, depth, seed=200 + i, mode=orc.FAITHFUL, threads=cores)
        paths += st.paths
        rays += st.segments
    dt = time.time() - t0
    value = paths / dt / 1e6
    sample = f"{w}x{h}, {step_spp} of {spp} spp per step, depth {depth} (paths/s is spp-independent)"
    line = {
        "impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": cfg, "scene": scene_name, "width": w, "height": h, "spp": spp, "max_depth": depth},  # same keys as the repo arm
        "sample": sample,
        "grays_per_s": rays / dt / 1e9,
        "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU restatement (oracle FAITHFUL mode) of the Rust reference, which cannot be built here (no cargo/rustc, rand not vendored)",
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(wl):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as orc

    scene_name, w, h, spp, depth, _ = wl
    cores = os.cpu_count() or 1
    s = orc.Scene(scene_name, 1)
    t0 = time.time()
    _, _, st = s.render(w, h, 1, depth, seed=1, mode=orc.FAITHFUL, threads=cores)
    rate = st.paths / max(time.time() - t0, 1e-6)
    sample_spp = max(1, min(spp, int(rate * 15.0 / (w * h))))
    t0 = time.time()
    _, _, st = s.render(w, h, sample_spp, depth, seed=2, mode=orc.FAITHFUL, threads=cores)
    dt = time.time() - t0
    return {"value": st.paths / dt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
            "sample": f"{w}x{h}, {sample_spp} of {spp} spp, depth {depth}, {dt:.1f} s of CPU time on {cores} threads",
            "grays_per_s": st.segments / dt / 1e9, "box_tests_per_ray": st.box_tests / st.segments, "prim_tests_per_ray": st.prim_tests / st.segments}


STAGES = ["generate", "extend", "shade", "accumulate"]


def measure(ctx, workload, steps, warmup, want_work=True):
    """Times `steps` renders of one workload on this rank's GPU (sharded over the ranks when world > 1) and returns the pieces of
    the JSON line that belong to it.  Collective: every rank calls it with the same arguments."""
    import numpy as np
    import torch
    import torch.distributed as dist

    rt, rtdist, dev, world, rank, args = ctx["rt"], ctx["rtdist"], ctx["dev"], ctx["world"], ctx["rank"], ctx["args"]
    scene_name, W, H, spp, depth, cfg = WORKLOADS[workload]
    flat = rt.FlatScene(scene_name, 1)
    scene = rt.Scene(dev, flat=flat)
    cam = flat.camera(W / H)
    stream = torch.cuda.current_stream()
    flush = ctx["flush"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(seed, flags=0):
        flush.zero_()
        return rtdist.render_sharded(scene, cam, W, H, spp, depth, seed, rank, world, args.queue_capacity, flags)

    sampler = ClockSampler(ctx["local_rank"])
    if rank == 0:
        sampler.start()  # before the warm-up: nvidia-smi needs a few hundred ms to come up
    for i in range(warmup):
        step(1000 + i)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_region0 = time.perf_counter()
    ev0.record(stream)
    paths = rays = launches = 0
    for i in range(steps):
        full, st = step(2000 + i)
        paths += st.paths
        rays += st.rays
        launches += st.launches + (1 if world > 1 else 0) + 1  # + de-interleave kernel, + the L2-flush memset
    ev1.record(stream)
    barrier()
    t_region1 = time.perf_counter()
    ms = ev0.elapsed_time(ev1)
    # ---- per-stage launch durations: the same steps again with CUDA events around the launches of every 8th wavefront iteration.
    # A stage-timed render runs as ONE wavefront lane, so that a timed launch has the device to itself (the timed region above runs
    # two interleaved lanes whose kernels share the SMs); these are the durations the rooflines are computed from. ----
    iters = 0
    ms_stage = np.zeros(4)
    for i in range(steps):
        _, st = step(2000 + i, rt.RENDER_STAGE_TIMING)
        iters += st.iterations
        ms_stage += [st.ms_generate, st.ms_extend, st.ms_shade, st.ms_accumulate]
    barrier()
    clocks = sampler.stop(t_region0, t_region1, time.perf_counter()) if rank == 0 else None
    red = torch.tensor([ms, *ms_stage], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(paths), float(rays), float(launches), float(iters)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_all = float(red[0])
    stage_max = [float(v) for v in red[1:]]  # slowest rank, summed over the steps
    tot_paths, tot_rays, tot_launches, tot_iters = float(tot[0]), float(tot[1]), int(tot[2]), float(tot[3])
    value = tot_paths / (ms_all * 1e-3) / 1e6
    mean_check = float(full.mean()) / spp

    # ---- algorithmic work per ray: one instrumented render outside the timed region (counts are spp-independent) ----
    box_per_ray = prim_per_ray = 0.0
    if want_work:
        _, stw = rtdist.render_sharded(scene, cam, W, H, min(spp, 16), depth, 4000, rank, world, args.queue_capacity, rt.RENDER_COUNT_WORK)
        work = torch.tensor([float(stw.box_tests), float(stw.prim_tests), float(stw.rays)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(work, op=dist.ReduceOp.SUM)
        box_per_ray, prim_per_ray = float(work[0] / work[2]), float(work[1] / work[2])

    # ---- e2e: host tables in, host framebuffer out, every step ------------------------------------
    scene_bytes = scene.device_bytes()
    if world == 1:
        host_fb = np.zeros((H, W, 3), dtype=np.float32)

        def e2e_step(seed):
            sc = rt.Scene(dev, flat=flat)  # rt_scene_create: BVH build + H2D of every table
            sc.render(cam, W, H, spp, depth, seed=seed, queue_capacity=args.queue_capacity, want_sumsq=False, out=host_fb)  # rt_render: host pointer out
            sc.close()
    else:
        host_fb = torch.empty((H, W, 3), dtype=torch.float32, pin_memory=True)

        def e2e_step(seed):
            sc = rt.Scene(dev, flat=flat)
            out, _ = rtdist.render_sharded(sc, cam, W, H, spp, depth, seed, rank, world, args.queue_capacity)
            if rank == 0:
                host_fb.copy_(out, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            sc.close()

    e2e_step(3000)
    barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        e2e_step(3001 + i)
    barrier()
    e2e_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = (W * H * spp * steps) / float(e2e_t[0]) / 1e6

    out = None
    if rank == 0:
        hbm_peak, peak_src, sm_max = load_peaks()
        info, fr = scene.bvh_info(), scene.frame_rect_info()
        kinds = flat.prims()["kind"]
        rect_frac = float(np.mean(kinds == rt.PRIM_RECT))
        extend_kernel = "k_extend_rects" if fr["frames"] else ("k_extend_bvh" if info["nodes"] >= 65536 else
                                                                ("k_extend<false, true>" if 0 < info["nodes"] <= 2048 else "k_extend<false, false>"))
        shade_kernel = "k_shade<2>" if fr["frames"] else "k_shade<1>"
        # small all-rectangle scenes: extend runs INSIDE k_shade (k_shade<1, false, true>: frame_rects_closest + shading per thread) —
        # no extend launch, no hit record, the ray read once
        fused = bool(fr["frames"]) and os.environ.get("RT_B200_FUSE_RECTS", "1") != "0"
        if fused:
            extend_kernel = shade_kernel = "k_shade<1, false, true>"
        kernel_of = {"generate": "k_generate", "extend": extend_kernel, "shade": shade_kernel, "accumulate": "k_accumulate"}
        launches_rank = tot_iters / world  # launches of each stage kernel on one rank = its wavefront iterations
        rays_rank = tot_rays / world
        rays_per_launch = tot_rays / max(tot_iters, 1)
        survive = 1.0 - tot_paths / tot_rays  # every path ends exactly once: the other segments scatter on
        # algorithmic HBM bytes per ray segment of the two queue-streaming stages: extend reads a, b (32 B) and writes
        # the hit (8 B); shade reads a, b, c and the hit (56 B) and writes a, b, c of the rays that scatter on (48 B each)
        # fused form: the one kernel reads a, b, c (48 B) and writes the survivors (48 B each); no hit record either way
        bytes_of = {"extend": float(EXTEND_BYTES_PER_RAY), "shade": (48.0 if fused else 56.0) + 48.0 * survive}
        dominant = max(("extend", "shade"), key=lambda s_: stage_max[STAGES.index(s_)])
        sm_mhz = (clocks or {}).get("sm_mhz") or sm_max

        def hbm_roofline(stage):
            dur = stage_max[STAGES.index(stage)] * 1e-3 / max(launches_rank, 1)
            achieved = bytes_of[stage] * rays_per_launch / max(dur, 1e-12) / 1e9
            bpr, src = ncu_traffic(workload, kernel_of[stage])
            return {"kernel": kernel_of[stage], "stage": stage, "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                    "traffic": bpr * rays_per_launch if bpr else None, "traffic_source": src, "traffic_measured_in_run": False,
                    "peak_source": peak_src, "launch_us": dur * 1e6, "rays_per_launch": rays_per_launch, "algorithmic_bytes_per_ray": bytes_of[stage],
                    "timing": "CUDA events around the kernel's launches of every 8th wavefront iteration, in a single-lane repeat of the timed "
                              "steps right after the timed region (the timed region itself overlaps two wavefront lanes)"}

        paths_rank = W * ((H + world - 1) // world) * spp
        huge = not fr["frames"] and paths_rank >= 1 << 31  # capi.cu render_lanes: 8 Mi-ray queues
        big = not fr["frames"] and paths_rank >= 1 << 29   # ... four lanes
        out = {
            "value": value, "ms_per_step": ms_all / steps, "steps": steps, "warmup": warmup,
            "config": {"workload": cfg, "scene": scene_name, "width": W, "height": H, "spp": spp, "max_depth": depth},
            "run": {"seed_scene": 1, "sharding": f"rows y % {world} == rank, NCCL all-gather + de-interleave" if world > 1 else "single GPU",
                    "l2": "flushed between steps (256 MiB memset inside the timed region)", "bvh": info, "frame_rects": fr,
                    "queue_capacity": args.queue_capacity or ((1 << 23) if huge else (1 << 22)),
                    "wavefront_lanes": int(os.environ.get("RT_B200_LANES", "3" if fr["frames"] else ("4" if big else "2")))},
            "grays_per_s": tot_rays / (ms_all * 1e-3) / 1e9, "rays_per_path": tot_rays / tot_paths, "mean_radiance": mean_check,
            "gpu_launches": tot_launches,
            "stage_ms_per_step": {s_: stage_max[i] / steps for i, s_ in enumerate(STAGES)},
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": scene_bytes + 64 + 40, "d2h_bytes_per_step": W * H * 3 * 4,
                    "ms_per_step": float(e2e_t[0]) / steps * 1e3,
                    "api": "rt_scene_create + rt_render (host pointers)" if world == 1 else "rt_scene_create + rt_render_device + NCCL all-gather + D2H on rank 0"},
            "roofline": hbm_roofline(dominant),
            "clocks": clocks,
        }
        if fused:
            out["roofline"]["fused"] = "extend + shade in one kernel: intersection (frame-rect list in shared memory) and shading per thread"
        else:
            other = "shade" if dominant == "extend" else "extend"
            out["roofline_" + other + "_hbm"] = hbm_roofline(other)
        if want_work:
            extend_s = stage_max[2 if fused else 1] * 1e-3  # fused: the intersection's instructions issue inside the shade stage
            instr_per_ray = box_per_ray * FP32_PER_BOX + prim_per_ray * (rect_frac * FP32_PER_RECT + (1 - rect_frac) * FP32_PER_SPHERE)
            rays_per_s_kernel = rays_rank / max(extend_s, 1e-12)  # per GPU, inside the extend kernel
            fp32_peak = 148 * 128 * sm_mhz * 1e6
            l2_bytes_per_ray = box_per_ray * 32 + prim_per_ray * (rect_frac * 64 + (1 - rect_frac) * 16)
            out["roofline_fp32"] = {"kernel": extend_kernel, "bound": "fp32-issue", "box_tests_per_ray": box_per_ray, "prim_tests_per_ray": prim_per_ray,
                                    "algorithmic_lane_instr_per_ray": instr_per_ray, "grays_per_s_in_kernel": rays_per_s_kernel / 1e9,
                                    "achieved": rays_per_s_kernel * instr_per_ray / 1e12, "peak": fp32_peak / 1e12, "unit": "T lane-instr/s",
                                    "frac": rays_per_s_kernel * instr_per_ray / fp32_peak, "sm_mhz": sm_mhz, "l2_bytes_per_ray": l2_bytes_per_ray,
                                    "note": "work counted by the instrumented GENERAL traversal on the scene's tables (RT_RENDER_COUNT_WORK), SURVEY §8(d) unit costs"}
            l2_peak = ctx.get("l2_peak")
            tables_in_smem = info["nodes"] <= 2048
            if l2_peak and not tables_in_smem:
                l2_gbs = l2_peak["l2_read_gbs_32MB"]
                out["roofline_l2"] = {"kernel": extend_kernel, "bound": "l2", "achieved": rays_per_s_kernel * l2_bytes_per_ray / 1e9, "peak": l2_gbs, "unit": "GB/s",
                                      "frac": rays_per_s_kernel * l2_bytes_per_ray / 1e9 / l2_gbs, "bytes_per_ray": l2_bytes_per_ray,
                                      "peak_source": "measured now by ray-tracer_b200/tools/l2bw (32 MB read stream, ld.global.cg)", "l2bw": l2_peak}
            elif l2_peak:
                out["roofline_l2"] = {"kernel": extend_kernel, "bound": "l2", "achieved": None, "peak": l2_peak["l2_read_gbs_32MB"], "unit": "GB/s", "frac": None,
                                      "l2bw": l2_peak, "note": "not applicable: this scene's node / primitive tables are staged in shared memory"}
        # the whole wavefront against HBM: every stage's algorithmic queue traffic over the timed step (lanes on).  Per ray segment:
        # extend 40 B, shade 56 B read; a surviving ray is written back (48 B), a finished path writes and accumulate reads a 16-byte
        # record, generate writes 48 B per path  ->  144 B x rays + 32 B x paths
        # (fused extend + shade: no extend pass over the rays and no hit record -> 96 B x rays + 32 B x paths)
        wf_bytes = (96.0 if fused else 144.0) * tot_rays + 32.0 * tot_paths
        wf_gbs = wf_bytes / (ms_all * 1e-3) / 1e9 / world  # per GPU
        out["roofline_wavefront_hbm"] = {"bound": "hbm", "achieved": wf_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": wf_gbs / hbm_peak,
                                         "bytes_per_ray": wf_bytes / max(tot_rays, 1.0), "per": "GPU, all four stage kernels over the timed step",
                                         "peak_source": peak_src}
    scene.close()
    return out


def measure_multi_c_abi(ctx, workload, steps):
    """rank 0: the same image through rt_render_multi (ONE process driving every GPU through the C ABI; host framebuffer out)."""
    rt, world = ctx["rt"], ctx["world"]
    scene_name, W, H, spp, depth, cfg = WORKLOADS[workload]
    flat = rt.FlatScene(scene_name, 1)
    multi = rt.MultiDevice(list(range(world)))
    ms = rt.MultiScene(multi, flat=flat)
    cam = flat.camera(W / H)
    import numpy as np
    buf = np.zeros((H, W, 3), dtype=np.float32)
    ms.render(cam, W, H, spp, depth, seed=1, want_sumsq=False, out=buf)
    ms.render(cam, W, H, spp, depth, seed=2, want_sumsq=False, out=buf)
    dev_ms, t0 = 0.0, time.perf_counter()
    for i in range(steps):
        _, _, st = ms.render(cam, W, H, spp, depth, seed=10 + i, want_sumsq=False, out=buf)
        dev_ms += st.ms_total
    wall = time.perf_counter() - t0
    paths = float(W) * H * spp * steps
    res = {"api": "rt_init_multi + rt_multi_scene_create + rt_render_multi (one process, one host thread per device, ncclCommInitAll, ncclSend/ncclRecv gather "
                  "to device 0, k_deinterleave, D2H)", "workload": cfg, "n_gpus": world, "steps": steps,
           "value": paths / (dev_ms * 1e-3) / 1e6, "unit": "Mpaths/s", "ms_per_step": dev_ms / steps,
           "e2e": {"value": paths / wall / 1e6, "unit": "Mpaths/s", "ms_per_step": wall / steps * 1e3, "d2h_bytes_per_step": W * H * 12},
           "mean_radiance": float(buf.mean()) / spp}
    ms.close()
    multi.close()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cornell-box", choices=sorted(WORKLOADS))
    ap.add_argument("--queue-capacity", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-book-one-4k", action="store_true", help="skip the north-star block (BASELINE.json configs[4]) of the line")
    ap.add_argument("--no-multi-c-abi", action="store_true", help="N > 1: skip the rt_render_multi (single process) block")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
        return

    import torch
    import torch.distributed as dist

    rt = importlib.import_module("ray-tracer_b200")  # raises if librt_b200.so is missing: there is no fallback
    rtdist = importlib.import_module("ray-tracer_b200.dist")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local_rank)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
        cpu_group = dist.new_group(backend="gloo")  # host-side barrier for the single-process phase (an NCCL barrier would spin ON the GPUs)
    ctx = {"rt": rt, "rtdist": rtdist, "dev": rt.Device(local_rank), "world": world, "rank": rank, "local_rank": local_rank, "args": args,
           "flush": torch.empty(256 << 20, dtype=torch.uint8, device="cuda"),  # > 126 MB L2
           "l2_peak": measure_l2() if rank == 0 else None}

    main_res = measure(ctx, args.workload, args.steps, args.warmup)
    # the north-star scene in the same run: fewer steps (1.6 s each at N = 1), same method
    b4k = None
    if args.workload != "book-one-4k" and not args.no_book_one_4k:
        b4k = measure(ctx, "book-one-4k", max(2, min(args.steps, 3)), 3)
    multi = None
    if world > 1 and not args.no_multi_c_abi:
        # the other ranks must leave their GPUs IDLE while rank 0 drives all of them from one process: they wait on the host (gloo);
        # a NCCL barrier would park a spinning kernel on every GPU and time-slice against rank 0's contexts
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
        if rank == 0:
            try:
                multi = {"headline": measure_multi_c_abi(ctx, args.workload, args.steps)}
                if b4k is not None:
                    multi["book_one_4k"] = measure_multi_c_abi(ctx, "book-one-4k", 2)
            except Exception as e:  # noqa: BLE001 - reported in the line, the per-rank measurement above stands on its own
                multi = {"error": str(e)}
        dist.barrier(group=cpu_group)

    if rank == 0:
        line = {"metric": "Mpaths/s", "value": main_res["value"], "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": main_res["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
        for k, v in main_res.items():
            if k not in ("value", "ms_per_step", "steps", "warmup"):
                line[k] = v
        if b4k is not None:
            b4k["metric"], b4k["unit"], b4k["n_gpus"] = "Mpaths/s", "Mpaths/s", world
            b4k["warp_execution_efficiency"] = {"kernel": "k_extend<false, true>", "value": 0.68, "measured_in_run": False,
                                                "source": "profiles/r2_50_full_metrics_book-one_extend_steady_state.txt: smsp__thread_inst_executed_per_inst_executed "
                                                          "21.7 / 32, mean of six consecutive steady-state launches (ncu --set full; hardware counters need the profiler)"}
            line["book_one_4k"] = b4k
        if multi is not None:
            line["multi_c_abi"] = multi
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
