#!/usr/bin/env python
"""bench.py — headline benchmark of the ray-tracing hot path on B200 (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c1|c3|c4|c5] [--impl ours|reference]

Default workload "c2" = BASELINE.json configs[1]: bunny-scale mesh (procedural stand-in, 81 920 triangles; the real
bunny.obj is missing from the checkout), 1920x1080, 1 spp, primary closest-hit rays (Camera::GenerateJitteredRay +
Scene::intersect single-ray rules) + one occlusion ray per hit toward the light (built as path_tracer.cpp:236-238).
metric = Mrays/s over both ray classes.  A "step" is one pass over the frame's rays.

  value  : rays resident in HBM when the timed region starts (prt_cuda_intersect_dev + prt_cuda_occluded_dev),
           CUDA events on the context's stream, L2 flushed between steps (inputs are smaller than L2).
  e2e    : the same pass through the host-buffer C ABI (prt_cuda_trace_jobs: the closest-hit batch and the occlusion
           batch in one call; the one-call-per-batch figure, prt_cuda_intersect + prt_cuda_occluded, is reported beside
           it as e2e.two_calls): pinned host rays in, H2D, kernels, D2H hit records / occlusion bytes out, all inside
           the timed region.
  N > 1  : every rank traces its own frame (different seed) — independent units, no data-path collective — "weak".
           Render workloads (c1/c3/c4/c5) instead shard the image rows of ONE frame across ranks and all-gather the
           framebuffer over NCCL ("strong").
  --impl reference : the UNMODIFIED reference (oracle/_ref/libprt_ref.so) on all host cores, same config/metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# bytes per ray of the roofline: 32 B per reference node pop + 48 B per leaf triangle + ray in / result out.
# Counter constants per workload are measured with the stats build of the reference / oracle and frozen in BASELINE.md.
TRAVERSAL_CONSTANTS = {
    # workload: (closest-hit nodes/ray, prims/ray, occlusion nodes/ray, prims/ray)
    "c2": (6.735, 3.594, 21.456, 8.062),
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the c2 kernels, from the `ncu --set full` capture of this
# command (profiles/r1/r1_g4_c2_trace_full_raw.csv): the tree is cache-resident, so DRAM traffic is far BELOW the
# algorithmic bytes (rays in + hits out + first touch of the tree).
NCU_TRAFFIC_C2 = {"intersect": 69.20e6 + 17.03e6, "occluded": 40.63e6 + 0.09e6}
# smsp__inst_executed.sum per launch of the same capture (2 073 600 closest-hit rays, 1 190 056 occlusion rays)
NCU_WARP_INST_C2 = {"intersect": 110324865.0, "occluded": 105084893.0}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel for render workloads")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = set()
        for r in self.rows:
            for i, nme in enumerate(names):
                if len(r) > 4 + i and r[4 + i].lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def c2_scene_and_rays(ctx, torch, seed, W=1920, H=1080):
    """Generates the frame's primary rays on the device (checked against the oracle in tests), traces them once to
    build the occlusion rays of the hits; returns device tensors (inputs resident in HBM)."""
    import prt_b200
    n = W * H
    d_rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    d_hits = torch.empty((n, 6), dtype=torch.float32, device="cuda")
    ctx.primary_dev((0, 0, W, H), 1, seed, prt_b200.PRT_MODE_SINGLE, d_rays.data_ptr(), None, d_hits.data_ptr())
    ctx.synchronize()
    return d_rays, d_hits


def shadow_rays_from(torch, d_rays, d_hits, light_dir, radius):
    """Occlusion rays exactly as path_tracer.cpp:236-238: org = pos + kFar*L, dir = -L, maxT = kFar - 0.0008 (fp32,
    unfused mul/add).  Input construction (outside the timed region)."""
    t = d_hits[:, 0]
    hit = t != -1.0
    r = d_rays[hit]
    tt = t[hit].unsqueeze(1)
    pos = tt * r[:, 3:6] + r[:, 0:3]
    L = torch.tensor(np.asarray(light_dir, np.float32), device="cuda")
    kfar = torch.tensor(np.float32(2.0) * np.float32(radius), device="cuda")
    sh = torch.zeros((pos.shape[0], 8), dtype=torch.float32, device="cuda")
    sh[:, 0:3] = pos + kfar * L
    sh[:, 3:6] = -L
    sh[:, 6] = kfar - torch.tensor(np.float32(0.0008), device="cuda")
    return sh.contiguous()


def north_star_primary(torch, local, flush_buf, steps=10):
    """BASELINE.json north_star target: primary-ray closest hit on Cornell+teapot, 1920x1080, 8 spp (one 8-ray packet
    per pixel, Camera::GenerateJitteredRayPacket + Scene::intersect<RayHitPacket,RayPacket> rules).  Rays resident in
    HBM, CUDA events on the context stream, L2 flushed between iterations.  Reported beside the headline, not as it."""
    import prt_b200
    from prt_b200 import scenes
    W, H = 1920, 1080
    ctx = prt_b200.load_scene(scenes.cornell_teapot(W, H, teapot_refl=scenes.SPECULAR, light=None), local)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=local)
    n = W * H * 8
    d_rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    d_avg = torch.empty((n // 8, 3), dtype=torch.float32, device="cuda")
    d_hits = torch.empty((n, 6), dtype=torch.float32, device="cuda")
    ctx.primary_dev((0, 0, W, H), 1, 1, prt_b200.PRT_MODE_PACKET, d_rays.data_ptr(), d_avg.data_ptr(), d_hits.data_ptr())
    ctx.synchronize()
    out = {}
    for name, fn in (("intersect_packet", lambda: ctx.intersect_dev(n, d_rays.data_ptr(), d_avg.data_ptr(), prt_b200.PRT_MODE_PACKET, d_hits.data_ptr())),
                     ("intersect_single", lambda: ctx.intersect_dev(n, d_rays.data_ptr(), None, prt_b200.PRT_MODE_SINGLE, d_hits.data_ptr())),
                     ("generate_and_intersect_packet", lambda: ctx.primary_dev((0, 0, W, H), 1, 1, prt_b200.PRT_MODE_PACKET, None, None, d_hits.data_ptr()))):
        ms = 0.0
        with torch.cuda.stream(ext):
            for k in range(3 + steps):
                flush_buf.zero_()
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record(ext); fn(); e1.record(ext)
                e1.synchronize()
                if k >= 3:
                    ms += e0.elapsed_time(e1)
        out[name] = {"mrays_per_s": n * steps / (ms * 1e-3) / 1e6, "ms": ms / steps}
    hitfrac = float((d_hits[:, 0] != -1.0).float().mean().item())
    ctx.close()
    return {"scene": "Cornell box + teapot 1920x1080, 8 primary rays per pixel", "rays": n, "hit_fraction": hitfrac, **out}


def dist_setup(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_reference(args, rank, world):
    """Reference arm: the unmodified reference on the host cores, same config/metric; rank 0 only."""
    if rank != 0:
        return
    from prt_b200 import scenes
    from oracle import refprt
    W, H = 1920, 1080
    threads = os.cpu_count() or 1
    kind = "reference" if refprt.available() else "port"
    if args.workload == "c2":
        desc = scenes.bunny_standin(W, H)
        sc = refprt.RefScene(desc) if kind == "reference" else refprt.OracleScene(desc)
        rays, hits = sc.primary_single((0, 0, W, H), 1, 1, nthreads=threads)
        L = np.asarray(desc["light"]["dir"], np.float32)
        kfar = np.float32(2.0) * np.float32(sc.radius())
        hit = hits[:, 0] != -1
        sh = np.zeros((int(hit.sum()), 8), np.float32)
        sh[:, 0:3] = (hits[hit, 0:1] * rays[hit, 3:6] + rays[hit, 0:3]) + kfar * L
        sh[:, 3:6] = -L
        sh[:, 6] = kfar - np.float32(0.0008)
        nrays = rays.shape[0] + sh.shape[0]
        with refprt.quiet_stdout():
            for _ in range(args.warmup):
                sc.intersect_single(rays, nthreads=threads); sc.occluded_single(sh, nthreads=threads)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                sc.intersect_single(rays, nthreads=threads); sc.occluded_single(sh, nthreads=threads)
            dt = time.perf_counter() - t0
        value = nrays * args.steps / dt / 1e6
        line = {"impl": "reference", "metric": "Mrays/s (closest-hit + occlusion)", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "c2: bunny-scale stand-in (81920 tris) 1920x1080 1spp primary closest-hit + occlusion rays, single-ray rules",
                           "rays_per_step": nrays},
                "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": f"{args.steps} full frames of {nrays} rays"},
                "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    else:
        cfg = render_config(args)
        desc = cfg["desc"]
        sc = refprt.RefScene(desc) if kind == "reference" else refprt.OracleScene(desc)
        # bounded sample: a centred band of rows sized for a few seconds per step
        rows = max(1, min(cfg["H"], int(cfg["ref_rows"])))
        y0 = (cfg["H"] - rows) // 2
        region = (0, y0, cfg["W"], y0 + rows)
        samples = cfg["W"] * rows * (cfg["spp"] // 8 * 8)
        with refprt.quiet_stdout():
            for _ in range(min(args.warmup, 1)):
                sc.render(cfg["spp"], 1, region=region, nthreads=threads)
            t0 = time.perf_counter()
            for s in range(args.steps):
                sc.render(cfg["spp"], 1 + s, region=region, nthreads=threads)
            dt = time.perf_counter() - t0
        value = samples * args.steps / dt / 1e6
        line = {"impl": "reference", "metric": "Msamples/s (full path tracing)", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": cfg["name"], "spp": cfg["spp"]},
                "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": threads, "kind": kind,
                                 "sample": f"rows {y0}..{y0 + rows} of {cfg['H']} ({samples} camera samples per step)"},
                "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def render_config(args):
    from prt_b200 import scenes
    wl = args.workload
    if wl == "c1":
        W, H, spp = 512, 512, 16
        desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
        name, ref_rows = "c1: Cornell box + teapot 512x512 16spp diffuse+emissive+directional", 512
    elif wl == "c3":
        W, H, spp = 1920, 1080, 256
        desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.SPECULAR, light="infinite")
        name, ref_rows = "c3: Cornell box + teapot 1920x1080 256spp full path tracing (specular teapot, infinite-area light)", 8
    elif wl == "c4":
        W, H, spp = 1920, 1080, 64
        desc = scenes.atrium(W, H)
        name, ref_rows = "c4: procedural Sponza-scale atrium (~262k tris, diffuse+alpha+bump textures) 1920x1080 64spp", 24
    elif wl == "c5":
        W, H, spp = 3840, 2160, 64
        desc = scenes.city(W, H)
        name, ref_rows = "c5: procedural San-Miguel-scale scene (~5M tris) 3840x2160 64spp", 16
    else:
        raise SystemExit(f"unknown workload {wl}")
    if args.spp:
        spp = args.spp
    return dict(W=W, H=H, spp=spp, desc=desc, name=name, ref_rows=ref_rows)


def run_ours(args, rank, world, local):
    import torch
    import prt_b200
    from prt_b200 import scenes
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hbm_peak, peak_src = peaks()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    flush_buf = torch.empty(384 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    if args.workload == "c2":
        W, H = 1920, 1080
        desc = scenes.bunny_standin(W, H)
        ctx = prt_b200.load_scene(desc, local)
        ext = torch.cuda.ExternalStream(ctx.stream(), device=local)
        d_rays, d_hits = c2_scene_and_rays(ctx, torch, 1 + rank)
        d_sh = shadow_rays_from(torch, d_rays, d_hits, desc["light"]["dir"], ctx.radius())
        n1, n2 = d_rays.shape[0], d_sh.shape[0]
        nrays = n1 + n2
        d_occ = torch.empty(n2, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()

        def step(ev=None):
            with torch.cuda.stream(ext):
                flush_buf.zero_()                                    # L2 flush between timed iterations (outside the events)
                if ev: ev[0].record(ext)
                ctx.intersect_dev(n1, d_rays.data_ptr(), None, prt_b200.PRT_MODE_SINGLE, d_hits.data_ptr())
                if ev: ev[1].record(ext)
                ctx.occluded_dev(n2, d_sh.data_ptr(), prt_b200.PRT_MODE_SINGLE, d_occ.data_ptr())
                if ev: ev[2].record(ext)

        for _ in range(max(args.warmup, 3)):
            step()
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        ctx.reset_counters()
        barrier()
        for k in range(args.steps):
            step(evs[k])
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        launches = ctx.counters()["kernelLaunches"]
        t_int = sum(e[0].elapsed_time(e[1]) for e in evs)
        t_occ = sum(e[1].elapsed_time(e[2]) for e in evs)
        total_ms = max_over_ranks(t_int + t_occ)
        all_rays = sum_over_ranks(float(nrays))
        value = all_rays * args.steps / (total_ms * 1e-3) / 1e6

        # ---- e2e: host buffers through prt_cuda_intersect / prt_cuda_occluded (H2D + kernels + D2H per step) ----
        h_rays = torch.empty((n1, 8), dtype=torch.float32).pin_memory(); h_rays.copy_(d_rays.cpu())
        h_sh = torch.empty((n2, 8), dtype=torch.float32).pin_memory(); h_sh.copy_(d_sh.cpu())
        h_hits = torch.empty((n1, 6), dtype=torch.float32).pin_memory()
        h_occ = torch.empty(n2, dtype=torch.uint8).pin_memory()
        L = ctx.L

        from prt_b200 import capi
        jobs = (capi.prt_trace_job * 2)()
        jobs[0].kind, jobs[0].n, jobs[0].mode, jobs[0].rays, jobs[0].out = prt_b200.PRT_JOB_INTERSECT, n1, prt_b200.PRT_MODE_SINGLE, h_rays.data_ptr(), h_hits.data_ptr()
        jobs[1].kind, jobs[1].n, jobs[1].mode, jobs[1].rays, jobs[1].out = prt_b200.PRT_JOB_OCCLUDED, n2, prt_b200.PRT_MODE_SINGLE, h_sh.data_ptr(), h_occ.data_ptr()

        def e2e_step_two_calls():                 # one synchronous call per batch: the copy pipeline fills and drains twice
            ctx.check(L.prt_cuda_intersect(ctx.h, n1, h_rays.data_ptr(), None, prt_b200.PRT_MODE_SINGLE, h_hits.data_ptr()))
            ctx.check(L.prt_cuda_occluded(ctx.h, n2, h_sh.data_ptr(), prt_b200.PRT_MODE_SINGLE, h_occ.data_ptr()))

        def e2e_step():                           # both batches handed over at once (prt_cuda_trace_jobs)
            ctx.check(L.prt_cuda_trace_jobs(ctx.h, 2, jobs))

        def time_e2e(f):
            for _ in range(3):
                f()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                f()
            barrier()
            return all_rays * args.steps / max_over_ranks(time.perf_counter() - t0) / 1e6

        e2e_two_calls = time_e2e(e2e_step_two_calls)
        occ_two_calls = int(h_occ.sum()); hits_two_calls = h_hits.clone()
        h_occ.zero_(); h_hits.zero_()
        e2e_value = time_e2e(e2e_step)
        assert int(h_occ.sum()) == int(d_occ.sum().item()) == occ_two_calls, "e2e and device-resident paths disagree"
        assert torch.equal(h_hits.view(torch.int32), hits_two_calls.view(torch.int32)) and torch.equal(h_hits.view(torch.int32), d_hits.cpu().view(torch.int32)), \
            "e2e hit records differ from the device-resident path"

        # secondary rooflines: the tree is cache-resident, so beside the HBM figure the kernel is placed against
        # the measured L2 read bandwidth and the measured FP32 warp-instruction issue rate of this device
        l2_peak, issue_peak = ctx.probe_peaks()
        cn, cp, on, op = TRAVERSAL_CONSTANTS["c2"]
        b_int = (32.0 * cn + 48.0 * cp + 56.0) * n1
        b_occ = (32.0 * on + 48.0 * op + 33.0) * n2
        dom = "occluded" if t_occ >= t_int else "intersect"
        dom_bytes, dom_ms = (b_occ, t_occ / args.steps) if dom == "occluded" else (b_int, t_int / args.steps)
        achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
        line = {"metric": "Mrays/s (closest-hit + occlusion)", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "c2: bunny-scale stand-in (81920 tris) 1920x1080 1spp primary closest-hit + occlusion rays, single-ray rules",
                           "rays_per_step_per_gpu": nrays, "closest_hit_rays": n1, "occlusion_rays": n2, "l2": "flushed between steps (384 MiB write)",
                           "per_rank": "own frame (seed 1+rank)"},
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 32 * nrays, "d2h_bytes_per_step": 24 * n1 + n2,
                        "call": "prt_cuda_trace_jobs (closest-hit batch + occlusion batch in one host-buffer call)",
                        "two_calls": {"value": e2e_two_calls, "call": "prt_cuda_intersect, then prt_cuda_occluded"}},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": f"{dom}Kernel<single>", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                             "frac": achieved / hbm_peak, "traffic": NCU_TRAFFIC_C2[dom], "algorithmic_bytes": dom_bytes,
                             "traffic_source": "ncu --set full, profiles/r1/r1_g4_c2_trace_full_raw.csv", "peak_source": peak_src,
                             "note": "algorithmic bytes (32 B/node pop + 48 B/leaf triangle + ray/result) / CUDA-event kernel time; the scene "
                                     "(nodes+triangles) is cache-resident, so this is NOT an HBM-limited kernel: frac > 1 means L1/L2 serve the re-reads",
                             "l2": {"peak": l2_peak, "unit": "GB/s", "frac": achieved / l2_peak,
                                    "peak_source": "prt_cuda_probe_peaks: 32 MiB L2-resident buffer, ld.global.cg.v4, all SMs"},
                             "issue": {"achieved": NCU_WARP_INST_C2[dom] / (dom_ms * 1e-3) / 1e9, "peak": issue_peak, "unit": "Gwarp-inst/s",
                                       "frac": NCU_WARP_INST_C2[dom] / (dom_ms * 1e-3) / 1e9 / issue_peak,
                                       "source": "smsp__inst_executed.sum per launch (profiles/r1/r1_g4_c2_trace_full_raw.csv) / live kernel time; "
                                                 "peak = prt_cuda_probe_peaks FFMA issue kernel"},
                             "kernel_ms": {"intersect": t_int / args.steps, "occluded": t_occ / args.steps},
                             "mrays_per_s": {"closest_hit": n1 / (t_int / args.steps) / 1e3, "occlusion": n2 / (t_occ / args.steps) / 1e3}},
                "clocks": clocks}
    else:
        line = run_render(args, rank, world, local, torch, dist, barrier, max_over_ranks, hbm_peak, peak_src)

    if rank == 0 and world == 1 and args.workload == "c2":
        line["north_star_primary"] = north_star_primary(torch, local, flush_buf)
    if rank == 0 and not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(args)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# per-workload traversal constants for the render workloads, measured with the oracle's counters on a sample region
# (tools/measure_constants.py): (extend nodes/ray, prims/ray, shadow nodes/ray, prims/ray, primary nodes/ray, prims/ray)
RENDER_CONSTANTS = {
    # scripts/measure_constants.py (oracle counters == the reference's PRT_ENABLE_STATS definitions), frozen in BASELINE.md §4
    "c1": (9.237, 19.565, 8.308, 9.830, 12.327, 14.534),
    "c3": (6.885, 17.347, 5.379, 8.187, 7.498, 9.272),
    "c4": (30.992, 27.685, 37.924, 12.986, 70.839, 16.089),
    "c5": (83.152, 65.959, 50.857, 17.493, 68.965, 12.170),
}


def run_render(args, rank, world, local, torch, dist, barrier, max_over_ranks, hbm_peak, peak_src):
    import prt_b200
    from prt_b200 import multigpu
    cfg = render_config(args)
    W, H, spp = cfg["W"], cfg["H"], cfg["spp"]
    ctx = prt_b200.load_scene(cfg["desc"], local)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=local)
    band = 8
    nb = (H + band - 1) // band
    nb_per_rank = (nb + world - 1) // world
    local_fb = torch.zeros((nb_per_rank * band, W, 3), dtype=torch.float32, device="cuda")
    gathered = torch.zeros((world, nb_per_rank * band, W, 3), dtype=torch.float32, device="cuda") if world > 1 else None
    host_fb = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    cam = cfg["desc"]["camera"]

    def step(seed, e2e=False):
        if e2e:
            ctx.set_camera(cam["pos"], cam["dir"], W, H)             # the per-step host input of a render is camera + seed
        ctx.render_bands_dev(band, rank, world, spp, seed, 1.0, local_fb.data_ptr())   # rank-local tile-major = all-gather send buffer
        with torch.cuda.stream(ext):
            ev = torch.cuda.Event()
            ev.record(ext)
        torch.cuda.current_stream().wait_event(ev)
        if world > 1:
            dist.all_gather_into_tensor(gathered.view(-1), local_fb.view(-1))     # the only collective: framebuffer gather over NVLink
            full = multigpu.detile(gathered.view(-1), world, band, W, H)
        else:
            full = local_fb[:H]
        if e2e:
            host_fb.copy_(full, non_blocking=True)
        return full

    for k in range(max(args.warmup, 3) if spp <= 64 else max(args.warmup, 1)):
        step(100 + k)
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ctx.reset_counters()
    barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    with torch.cuda.stream(ext):
        s0 = torch.cuda.Event(); s0.record(ext)        # order ctx-stream work after e0 is irrelevant for timing; total is bracketed on the default stream
    for k in range(args.steps):
        full = step(1 + k)
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = max_over_ranks(e0.elapsed_time(e1))
    cnt = ctx.counters()
    samples = float(W) * H * (spp // 8 * 8)
    value = samples * args.steps / (total_ms * 1e-3) / 1e6
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(1 + k, e2e=True)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    line = {"metric": "Msamples/s (full path tracing)", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": cfg["name"], "spp": spp, "sharding": f"{band}-row bands dealt round-robin to {world} rank(s), scene replicated, NCCL all-gather of the framebuffer",
                       "l2": "per-step working set (wavefront state + queues) exceeds L2"},
            "e2e": {"value": samples * args.steps / e2e_s / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": 56, "d2h_bytes_per_step": int(host_fb.numel() * 4)},
            "gpu_launches": int(cnt["kernelLaunches"]),
            "rays": {"primary": cnt["primaryRays"], "extend": cnt["extendRays"], "shadow": cnt["shadowRays"],
                     "shadow_not_traced_zero_radiance": cnt["shadowRaysSkipped"]},
            "roofline": None, "clocks": clocks}
    tot_rays = cnt["primaryRays"] + cnt["extendRays"] + cnt["shadowRays"]
    line["mrays_per_s"] = tot_rays * world / (total_ms * 1e-3) / 1e6 if cnt["extendRays"] else None
    k = RENDER_CONSTANTS.get(args.workload)
    if k and cnt["extendRays"]:
        b = (32 * k[0] + 48 * k[1] + 56) * cnt["extendRays"] + (32 * k[2] + 48 * k[3] + 33) * cnt["shadowRays"] + (32 * k[4] + 48 * k[5] + 56) * cnt["primaryRays"]
        ach = b / (total_ms * 1e-3) / 1e9
        line["roofline"] = {"bound": "hbm", "kernel": "wfTrace+wfPrimary", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                            "traffic": None, "peak_source": peak_src,
                            "note": "whole step; algorithmic bytes = sum over the step's rays of 32 B/node pop + 48 B/triangle + ray/result "
                                    "(per-class constants measured with the oracle's counters); > 1 means caches serve the re-reads"}
    return line


def cpu_baseline(args):
    """Reference (or oracle port) timed on this box's host cores on a bounded sample of the same workload."""
    from prt_b200 import scenes
    from oracle import refprt
    threads = os.cpu_count() or 1
    kind = "reference" if refprt.available() else "port"
    if kind == "port":
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    if args.workload == "c2":
        W, H = 1920, 1080
        desc = scenes.bunny_standin(W, H)
        with refprt.quiet_stdout():
            sc = refprt.RefScene(desc) if kind == "reference" else refprt.OracleScene(desc)
            rays, hits = sc.primary_single((0, 0, W, H), 1, 1, nthreads=threads)
            L = np.asarray(desc["light"]["dir"], np.float32)
            kfar = np.float32(2.0) * np.float32(sc.radius())
            hit = hits[:, 0] != -1
            sh = np.zeros((int(hit.sum()), 8), np.float32)
            sh[:, 0:3] = (hits[hit, 0:1] * rays[hit, 3:6] + rays[hit, 0:3]) + kfar * L
            sh[:, 3:6] = -L
            sh[:, 6] = kfar - np.float32(0.0008)
            sc.intersect_single(rays, nthreads=threads)
            reps, t0 = 0, time.perf_counter()
            while time.perf_counter() - t0 < 10.0:
                sc.intersect_single(rays, nthreads=threads); sc.occluded_single(sh, nthreads=threads); reps += 1
            dt = time.perf_counter() - t0
        n = rays.shape[0] + sh.shape[0]
        return {"value": n * reps / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
                "sample": f"{reps} full frames ({n} rays each: {rays.shape[0]} closest-hit + {sh.shape[0]} occlusion) in {dt:.1f} s"}
    cfg = render_config(args)
    with refprt.quiet_stdout():
        sc = refprt.RefScene(cfg["desc"]) if kind == "reference" else refprt.OracleScene(cfg["desc"])
        rows = max(1, min(cfg["H"], int(cfg["ref_rows"])))
        y0 = (cfg["H"] - rows) // 2
        t0 = time.perf_counter()
        sc.render(cfg["spp"], 1, region=(0, y0, cfg["W"], y0 + rows), nthreads=threads)
        dt = time.perf_counter() - t0
    samples = cfg["W"] * rows * (cfg["spp"] // 8 * 8)
    return {"value": samples / dt / 1e6, "unit": "Msamples/s", "cores": threads, "kind": kind,
            "sample": f"rows {y0}..{y0 + rows} of {cfg['H']} at {cfg['spp']} spp ({samples} camera samples) in {dt:.1f} s"}


def main():
    args = parse()
    rank, world, local = dist_setup(args)
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_ours(args, rank, world, local)


if __name__ == "__main__":
    main()
