#!/usr/bin/env python
"""bench.py — headline benchmark of the ray-tracing hot path on B200 (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload all|c2|c1|c3|c4|c5] [--impl ours|reference]

One JSON line.  Its top level is BASELINE.json's ray metric on configs[1] ("c2": bunny-scale mesh — procedural stand-in,
81 920 triangles, the real bunny.obj is missing from the checkout — 1920x1080, primary closest-hit rays + one occlusion ray
per hit toward the light, built as path_tracer.cpp:236-238), its `render` block the samples/s half of the metric on the
path-tracing configs (c1, c3, c4, c5 at N = 1; c3 and c5 — the ones BASELINE.json quotes at 1/2/4/8 GPUs — at N > 1).

  rules  : the reference traces primary rays through its 8-wide packet instantiation (Scene::intersect<RayHitPacket,RayPacket>,
           path_tracer.cpp:57-75: 8 jittered rays per pixel from Camera::GenerateJitteredRayPacket) and every other ray through
           the single-ray one.  Both rule sets are timed on both arms (`rules.packet`, `rules.single`); the headline `value` is
           the PACKET pair, the reference's faster stock path.
  value  : rays resident in HBM when the timed region starts (prt_cuda_intersect_dev + prt_cuda_occluded_dev), CUDA events
           on the context's stream, L2 flushed between passes.  A step = `passes_per_step` passes over the frame's rays, so that
           the K timed steps span about a second.
  e2e    : the same frame through the host-buffer C ABI, H2D + kernels + D2H inside the timed region.  `e2e.value` is the
           compact form (prt_cuda_trace_jobs with a PRT_RAYS_PRIMARY job and a PRT_RAYS_FROM_HITS job: camera + seed in,
           hit records + occlusion bytes out); `e2e.plain` sends the same rays as 32-byte records (round-1 path),
           `e2e.shared` as 12-byte directions / origins, `e2e.pageable` the plain path from unpinned numpy memory.
  render : PathTracer::TraceBlock over the whole frame (path_tracer.cpp:17-33, main.cpp:121-158) through prt_cuda_render_dev.
           At N > 1 the frame's rows are sharded over the N ranks in bands and the framebuffer is all-gathered with NCCL
           INSIDE libprt_cuda.so and inside the timed region (prt_cuda_comm_init_rank; torch.distributed only broadcasts the
           communicator id and takes max-over-ranks).  `image_crc` is the CRC-32 of the seed-1 frame: equal at every N.
  N > 1  : c2 has no exchange step: every rank traces its own frame (seed 1 + rank) — "weak".  Renders shard ONE frame — "strong".
  --impl reference : the UNMODIFIED reference (oracle/_ref/libprt_ref.so) on all host cores, same configs and metrics.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W2, H2 = 1920, 1080
C2_NAME = "c2: bunny-scale stand-in (81920 tris) 1920x1080 primary closest-hit + occlusion rays (headline: packet rules, 8 rays/pixel; single rules, 1 ray/pixel, beside it)"

# bytes per ray of the roofline: 32 B per reference node pop + 48 B per leaf triangle + ray in / result out.
# Counters per ray class measured with the reference's PRT_ENABLE_STATS build / the oracle's counters (scripts/measure_constants.py),
# frozen in BASELINE.md §4: (closest-hit nodes/ray, prims/ray, occlusion nodes/ray, prims/ray)
TRAVERSAL_CONSTANTS = {"c2_single": (6.735, 3.594, 21.456, 8.062), "c2_packet": (12.782, 3.498, 21.580, 8.127)}   # packet: oracle counters on every 9th row of the frame
# render workloads: (extend nodes/ray, prims/ray, shadow nodes/ray, prims/ray, primary nodes/ray, prims/ray)
RENDER_CONSTANTS = {
    "c1": (9.237, 19.565, 8.308, 9.830, 12.327, 14.534),
    "c3": (6.885, 17.347, 5.379, 8.187, 7.498, 9.272),
    "c4": (30.992, 27.685, 37.924, 12.986, 70.839, 16.089),
    "c5": (83.152, 65.959, 50.857, 17.493, 68.965, 12.170),
}
# ncu-derived per-launch numbers (dram bytes, warp instructions, L1 data-pipe wavefronts) of the dominant kernels: written by
# scripts/ncu_constants.py from the round's `ncu --set full` captures, together with the commit they were captured at.
NCU_CONSTANTS_FILE = os.path.join(ROOT, "profiles", "r2", "ncu_constants.json")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="all", help="all | c2 | c1 | c3 | c4 | c5 (one piece only, for profiling)")
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fetched", action="store_true", help="skip the counting-build pass (fetched bytes per ray)")
    ap.add_argument("--no-fast", action="store_true", help="skip the opt-in fast traversal mode's record (render.<wl>.fast_mode)")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel for render workloads")
    ap.add_argument("--render-steps", type=int, default=0, help="override the timed steps of each render workload")
    ap.add_argument("--passes", type=int, default=0, help="override c2 passes per step")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed regions (B200_PROFILING.md recipe)."""

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
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
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


def ncu_constants():
    if os.path.exists(NCU_CONSTANTS_FILE):
        with open(NCU_CONSTANTS_FILE) as f:
            return json.load(f)
    return {}


def c2_config(n1s, n2s, n1p, n2p, passes):
    """The config dict both arms print (identical by construction: every number is a property of the workload)."""
    return {"workload": C2_NAME, "rays_per_pass_per_gpu": n1p + n2p, "closest_hit_rays": n1p, "occlusion_rays": n2p,
            "single_rules": {"closest_hit_rays": n1s, "occlusion_rays": n2s}, "passes_per_step": passes,
            "l2": "flushed between passes (384 MiB write)", "per_rank": "own frame (seed 1 + rank)"}


def render_config(wl, spp_override=0):
    from prt_b200 import scenes
    if wl == "c1":
        W, H, spp = 512, 512, 16
        desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
        name, ref_tiles, steps = "c1: Cornell box + teapot 512x512 16spp diffuse+emissive+directional", 1024, 40
    elif wl == "c3":
        W, H, spp = 1920, 1080, 256
        desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.SPECULAR, light="infinite")
        name, ref_tiles, steps = "c3: Cornell box + teapot 1920x1080 256spp full path tracing (specular teapot, infinite-area light)", 128, 3
    elif wl == "c4":
        W, H, spp = 1920, 1080, 64
        desc = scenes.atrium(W, H)
        name, ref_tiles, steps = "c4: procedural Sponza-scale atrium (~262k tris, diffuse+alpha+bump textures) 1920x1080 64spp", 512, 5
    elif wl == "c5":
        W, H, spp = 3840, 2160, 64
        desc = scenes.city(W, H)
        name, ref_tiles, steps = "c5: procedural San-Miguel-scale scene (~5M tris) 3840x2160 64spp", 512, 2
    else:
        raise SystemExit(f"unknown workload {wl}")
    if spp_override:
        spp = spp_override
    return dict(wl=wl, W=W, H=H, spp=spp, desc=desc, name=name, ref_tiles=ref_tiles, steps=steps)


def stratified_tiles(W, H, count, tile=16):
    """`count` tiles of 16x16 pixels (the reference's own work unit, main.cpp:123-138) on a regular lattice over the frame's tile
    grid, so the CPU sample sees every part of the image the GPU renders and every host thread stays busy."""
    tx, ty = (W + tile - 1) // tile, (H + tile - 1) // tile
    total = tx * ty
    count = max(1, min(total, int(count)))
    ids = sorted(set(int((k + 0.5) * total / count) for k in range(count)))
    return [((i % tx) * tile, (i // tx) * tile, min(W, (i % tx + 1) * tile), min(H, (i // tx + 1) * tile)) for i in ids], total


def shadow_rays_np(rays, hits, L, radius):
    """path_tracer.cpp:236-238 in fp32 (unfused): org = (t*dir + org) + kFar*L, dir = -L, maxT = kFar - 0.0008."""
    L = np.asarray(L, np.float32)
    kfar = np.float32(2.0) * np.float32(radius)
    hit = hits[:, 0] != -1
    sh = np.zeros((int(hit.sum()), 8), np.float32)
    sh[:, 0:3] = (hits[hit, 0:1] * rays[hit, 3:6] + rays[hit, 0:3]) + kfar * L
    sh[:, 3:6] = -L
    sh[:, 6] = kfar - np.float32(0.0008)
    return sh


# ---------------------------------------------------------------------------------------------------------------------
# reference side (CPU): used by --impl reference and by the cpu_baseline legs of our arm (rank 0, N = 1)
# ---------------------------------------------------------------------------------------------------------------------
class RefC2:
    """The c2 frame on the reference: rays of both rule sets, their hit records, and timed passes."""

    def __init__(self):
        from prt_b200 import scenes
        from oracle import refprt
        self.refprt = refprt
        self.threads = os.cpu_count() or 1
        self.kind = "reference" if refprt.available() else "port"
        if self.kind == "port":
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])
        desc = scenes.bunny_standin(W2, H2)
        with refprt.quiet_stdout():
            self.sc = refprt.RefScene(desc) if self.kind == "reference" else refprt.OracleScene(desc)
            L = desc["light"]["dir"]
            t = self.threads
            self.s_rays, self.s_hits = self.sc.primary_single((0, 0, W2, H2), 1, 1, nthreads=t)
            self.s_sh = shadow_rays_np(self.s_rays, self.s_hits, L, self.sc.radius())
            self.p_rays, self.p_avg, self.p_hits = self.sc.primary_packets((0, 0, W2, H2), 1, 1, nthreads=t)
            sh = shadow_rays_np(self.p_rays, self.p_hits, L, self.sc.radius())
            self.p_sh = np.ascontiguousarray(sh[: sh.shape[0] // 8 * 8])            # whole packets of 8 shadow rays
            self.p_mask = np.full(self.p_sh.shape[0] // 8, 0xff, np.uint8)

    def counts(self):
        return self.s_rays.shape[0], self.s_sh.shape[0], self.p_rays.shape[0], self.p_sh.shape[0]

    def pass_single(self):
        self.sc.intersect_single(self.s_rays, nthreads=self.threads); self.sc.occluded_single(self.s_sh, nthreads=self.threads)

    def pass_packet(self):
        self.sc.intersect_packet(self.p_rays, self.p_avg, nthreads=self.threads); self.sc.occluded_packet(self.p_sh, self.p_mask, nthreads=self.threads)

    def timed(self, fn, nrays, budget_s=None, passes=None, warm=1):
        with self.refprt.quiet_stdout():
            for _ in range(warm):
                fn()
            reps, t0 = 0, time.perf_counter()
            while (passes is not None and reps < passes) or (passes is None and time.perf_counter() - t0 < budget_s):
                fn(); reps += 1
            dt = time.perf_counter() - t0
        return nrays * reps / dt / 1e6, reps, dt


def ref_render_sample(cfg, threads, kind, refprt, repeats=1):
    """Reference TraceBlock on 16x16 tiles spread evenly over the frame, one tile per task on all host cores (the reference's
    own driver loop, main.cpp:121-158); returns (Msamples/s, description)."""
    with refprt.quiet_stdout():
        sc = refprt.RefScene(cfg["desc"]) if kind == "reference" else refprt.OracleScene(cfg["desc"])
        tiles, total = stratified_tiles(cfg["W"], cfg["H"], cfg["ref_tiles"])
        out = np.zeros((cfg["H"], cfg["W"], 3), np.float32)
        t0 = time.perf_counter()
        for rep in range(repeats):
            sc.render_tiles(tiles, cfg["spp"], 1 + rep, nthreads=threads, out=out)
        dt = time.perf_counter() - t0
    pixels = sum((t[2] - t[0]) * (t[3] - t[1]) for t in tiles)
    samples = pixels * (cfg["spp"] // 8 * 8) * repeats
    what = "the whole frame" if len(tiles) == total else f"{len(tiles)} of the frame's {total} 16x16 tiles, spread evenly"
    return samples / dt / 1e6, f"{what}, one tile per task, at {cfg['spp']} spp x {repeats} ({samples} camera samples) in {dt:.1f} s"


def run_reference(args, rank, world):
    """Reference arm: the unmodified reference on the host cores, same config/metric; rank 0 only."""
    if rank != 0:
        return
    from oracle import refprt
    line = {"impl": "reference"}
    threads = os.cpu_count() or 1
    kind = "reference" if refprt.available() else "port"
    if args.workload in ("all", "c2"):
        ref = RefC2()
        n1s, n2s, n1p, n2p = ref.counts()
        passes = args.passes or 128
        # every step = a bounded sample of the workload: one pass over the frame's rays per rule set
        with refprt.quiet_stdout():
            for _ in range(min(args.warmup, 2)):
                ref.pass_packet()
        vp, reps_p, dt_p = ref.timed(ref.pass_packet, n1p + n2p, passes=args.steps, warm=0)
        vs, reps_s, dt_s = ref.timed(ref.pass_single, n1s + n2s, passes=max(2, args.steps // 4), warm=1)
        line.update({"metric": "Mrays/s (closest-hit + occlusion)", "value": vp, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
                     "warmup": args.warmup, "ms_per_step": dt_p / reps_p * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                     "dtype": "f32", "data": "synthetic", "config": c2_config(n1s, n2s, n1p, n2p, passes),
                     "rules": {"packet": {"value": vp, "unit": "Mrays/s", "rays_per_pass": n1p + n2p, "passes": reps_p,
                                          "call": "Scene::intersect<RayHitPacket,RayPacket> + Scene::occluded<RayPacketMask,RayPacket>"},
                               "single": {"value": vs, "unit": "Mrays/s", "rays_per_pass": n1s + n2s, "passes": reps_s,
                                          "call": "Scene::intersect<SingleRayHitPacket,SingleRayPacket> + Scene::occluded<bool,SingleRayPacket>"},
                               "headline": "packet"},
                     "cpu_baseline": {"value": vp, "unit": "Mrays/s", "cores": threads, "kind": kind,
                                      "sample": f"{reps_p} passes over the frame's {n1p + n2p} packet-rule rays (one pass per step; our arm runs {passes} passes per step)"},
                     "e2e": {"value": vp, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    # the same render workloads as our arm at this N (all four at N = 1; the configs quoted at 1/2/4/8 GPUs, C3 and C5, at N > 1)
    if args.workload == "all":
        wls = ["c1", "c3", "c4", "c5"] if args.gpus == 1 else ["c3", "c5"]
    else:
        wls = [] if args.workload == "c2" else [args.workload]
    render = {}
    for wl in wls:
        cfg = render_config(wl, args.spp)
        v, what = ref_render_sample(cfg, threads, kind, refprt, repeats=3 if wl == "c1" else 1)
        render[wl] = {"value": v, "unit": "Msamples/s", "config": {"workload": cfg["name"], "spp": cfg["spp"]},
                      "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": threads, "kind": kind, "sample": what}}
    if render:
        line["render"] = render
    if "metric" not in line:                      # a single render workload was asked for: it is the line's metric
        wl = wls[0]
        cfg = render_config(wl, args.spp)
        line.update({"metric": "Msamples/s (full path tracing)", "value": render[wl]["value"], "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
                     "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                     "data": "synthetic", "config": {"workload": cfg["name"], "spp": cfg["spp"]}, "cpu_baseline": render[wl]["cpu_baseline"],
                     "e2e": {"value": render[wl]["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# our side
# ---------------------------------------------------------------------------------------------------------------------
class Env:
    def __init__(self, args):
        import torch
        self.torch = torch
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        self.hbm_peak, self.peak_src = peaks()
        self.flush_buf = torch.empty(384 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
        self.launches = 0
        self.comm_id = None

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return float(t.item())

    def join_comm(self, ctx):
        """prt_cuda_comm_init_rank on this rank's context: torch.distributed only carries the 128-byte NCCL id."""
        if self.world <= 1:
            return
        from prt_b200 import capi
        torch = self.torch
        ident = torch.zeros(capi.PRT_COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
        if self.rank == 0:
            ident = torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8).cuda()
        self.dist.broadcast(ident, 0)
        ctx.comm_init_rank(ident.cpu().numpy().tobytes(), self.rank, self.world)


def run_c2(env, sampler_clocks):
    import prt_b200
    from prt_b200 import capi, scenes
    torch, args, rank, world, local = env.torch, env.args, env.rank, env.world, env.local
    SINGLE, PACKET = prt_b200.PRT_MODE_SINGLE, prt_b200.PRT_MODE_PACKET
    desc = scenes.bunny_standin(W2, H2)
    ctx = prt_b200.load_scene(desc, local)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=local)
    seed = 1 + rank
    Ldir = np.asarray(desc["light"]["dir"], np.float32)
    radius = ctx.radius()

    def shadow_rays_dev(d_rays, d_hits, whole_packets):
        t = d_hits[:, 0]
        hit = t != -1.0
        r = d_rays[hit]
        pos = t[hit].unsqueeze(1) * r[:, 3:6] + r[:, 0:3]
        L = torch.tensor(Ldir, device="cuda")
        kfar = torch.tensor(np.float32(2.0) * np.float32(radius), device="cuda")
        sh = torch.zeros((pos.shape[0], 8), dtype=torch.float32, device="cuda")
        sh[:, 0:3] = pos + kfar * L
        sh[:, 3:6] = -L
        sh[:, 6] = kfar - torch.tensor(np.float32(0.0008), device="cuda")
        if whole_packets:
            sh = sh[: sh.shape[0] // 8 * 8]
        return sh.contiguous()

    # ---- inputs resident in HBM: both rule sets' rays, generated on the device (checked against the oracle in tests) ----
    sets = {}
    for name, mode, per in (("packet", PACKET, 8), ("single", SINGLE, 1)):
        n1 = W2 * H2 * per
        d_rays = torch.empty((n1, 8), dtype=torch.float32, device="cuda")
        d_avg = torch.empty((n1 // 8, 3), dtype=torch.float32, device="cuda") if mode == PACKET else None
        d_hits = torch.empty((n1, 6), dtype=torch.float32, device="cuda")
        ctx.primary_dev((0, 0, W2, H2), 1, seed, mode, d_rays.data_ptr(), d_avg.data_ptr() if d_avg is not None else None, d_hits.data_ptr())
        ctx.synchronize()
        d_sh = shadow_rays_dev(d_rays, d_hits, mode == PACKET)
        d_occ = torch.empty(d_sh.shape[0], dtype=torch.uint8, device="cuda")
        sets[name] = dict(mode=mode, n1=n1, n2=d_sh.shape[0], d_rays=d_rays, d_avg=d_avg, d_hits=d_hits, d_sh=d_sh, d_occ=d_occ)
    torch.cuda.synchronize()

    def one_pass(S, ev=None):
        with torch.cuda.stream(ext):
            env.flush_buf.zero_()                                    # L2 flush between timed passes (outside the events)
            if ev: ev[0].record(ext)
            ctx.intersect_dev(S["n1"], S["d_rays"].data_ptr(), S["d_avg"].data_ptr() if S["d_avg"] is not None else None, S["mode"], S["d_hits"].data_ptr())
            if ev: ev[1].record(ext)
            ctx.occluded_dev(S["n2"], S["d_sh"].data_ptr(), S["mode"], S["d_occ"].data_ptr())
            if ev: ev[2].record(ext)

    def timed_passes(S, steps, passes):
        for _ in range(max(args.warmup, 3)):
            one_pass(S)
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps * passes)]
        ctx.reset_counters()
        env.barrier()
        for e in evs:
            one_pass(S, e)
        env.barrier()
        env.launches += int(ctx.counters()["kernelLaunches"])
        t_int = sum(e[0].elapsed_time(e[1]) for e in evs)
        t_occ = sum(e[1].elapsed_time(e[2]) for e in evs)
        total_ms = env.reduce(t_int + t_occ, "MAX")
        all_rays = env.reduce(float(S["n1"] + S["n2"]), "SUM")
        return dict(value=all_rays * steps * passes / (total_ms * 1e-3) / 1e6, total_ms=total_ms, t_int=t_int / (steps * passes), t_occ=t_occ / (steps * passes))

    # passes per step: sized so that the K timed steps span about a second (r1's 8.5 ms region could not show sustained clocks)
    passes = args.passes or 128
    P, S = sets["packet"], sets["single"]
    rp = timed_passes(P, args.steps, max(1, passes // 4))           # a packet pass is ~4x a single pass (8 rays per pixel)
    rs = timed_passes(S, args.steps, passes)

    # ---- e2e: host buffers through prt_cuda_trace_jobs (H2D + kernels + D2H per step) ----
    L = ctx.L
    region = (0, 0, W2, H2)

    def make_jobs(specs):
        jobs = (capi.prt_trace_job * len(specs))()
        for J, sp in zip(jobs, specs):
            for k, v in sp.items():
                if k in ("shared", "region"):
                    getattr(J, k)[:] = v
                else:
                    setattr(J, k, v)
        return jobs

    def time_e2e(jobs, nrays, reps):
        for _ in range(2):
            ctx.check(L.prt_cuda_trace_jobs(ctx.h, len(jobs), jobs))
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            ctx.check(L.prt_cuda_trace_jobs(ctx.h, len(jobs), jobs))
        env.barrier()
        return env.reduce(float(nrays), "SUM") * reps / env.reduce(time.perf_counter() - t0, "MAX") / 1e6

    e2e = {}
    for name, Sx in (("packet", P), ("single", S)):
        n1, n2, mode = Sx["n1"], Sx["n2"], Sx["mode"]
        pin = lambda shape, dt: torch.empty(shape, dtype=dt).pin_memory()
        h_rays = pin((n1, 8), torch.float32); h_rays.copy_(Sx["d_rays"])
        h_sh = pin((n2, 8), torch.float32); h_sh.copy_(Sx["d_sh"])
        h_avg = None
        if mode == PACKET:
            h_avg = pin((n1 // 8, 3), torch.float32); h_avg.copy_(Sx["d_avg"])
        h_hits = pin((n1, 6), torch.float32); h_occ = pin((n2,), torch.uint8); h_occ1 = pin((n1,), torch.uint8)
        h_dirs = pin((n1, 3), torch.float32); h_dirs.copy_(Sx["d_rays"][:, 3:6])
        h_org = pin((n2, 3), torch.float32); h_org.copy_(Sx["d_sh"][:, 0:3])
        cam = [float(x) for x in desc["camera"]["pos"]]
        sh0 = Sx["d_sh"][0].cpu().numpy()
        plain = make_jobs([dict(kind=0, n=n1, mode=mode, layout=capi.PRT_RAYS_FULL, rays=h_rays.data_ptr(), avgDir=h_avg.data_ptr() if h_avg is not None else None, out=h_hits.data_ptr()),
                           dict(kind=1, n=n2, mode=mode, layout=capi.PRT_RAYS_FULL, rays=h_sh.data_ptr(), out=h_occ.data_ptr())])
        shared = make_jobs([dict(kind=0, n=n1, mode=mode, layout=capi.PRT_RAYS_SHARED_ORIGIN, rays=h_dirs.data_ptr(), avgDir=h_avg.data_ptr() if h_avg is not None else None,
                                 out=h_hits.data_ptr(), shared=cam + [100000.0]),
                            dict(kind=1, n=n2, mode=mode, layout=capi.PRT_RAYS_SHARED_DIR, rays=h_org.data_ptr(), out=h_occ.data_ptr(),
                                 shared=[float(sh0[3]), float(sh0[4]), float(sh0[5]), float(sh0[6])])])
        compact = make_jobs([dict(kind=0, n=n1, mode=mode, layout=capi.PRT_RAYS_PRIMARY, out=h_hits.data_ptr(), region=list(region), perPixel=1, seed=seed),
                             dict(kind=1, n=n1, mode=mode, layout=capi.PRT_RAYS_FROM_HITS, out=h_occ1.data_ptr(), fromJob=0, shared=[float(x) for x in Ldir] + [0.0])])
        reps = max(3, args.steps // (4 if mode == PACKET else 1))
        r = {"plain": {"value": time_e2e(plain, n1 + n2, reps), "h2d_bytes": 32 * (n1 + n2) + (12 * (n1 // 8) if mode == PACKET else 0), "d2h_bytes": 24 * n1 + n2,
                       "call": "prt_cuda_trace_jobs, PRT_RAYS_FULL x 2 (32-byte rays up)"}}
        hits_plain = h_hits.clone(); occ_plain = h_occ.clone(); h_hits.zero_(); h_occ.zero_()
        r["shared"] = {"value": time_e2e(shared, n1 + n2, reps), "h2d_bytes": 12 * (n1 + n2) + (12 * (n1 // 8) if mode == PACKET else 0), "d2h_bytes": 24 * n1 + n2,
                       "call": "prt_cuda_trace_jobs, PRT_RAYS_SHARED_ORIGIN + PRT_RAYS_SHARED_DIR (12 bytes per ray up)"}
        assert torch.equal(h_hits.view(torch.int32), hits_plain.view(torch.int32)) and torch.equal(h_occ, occ_plain), "shared-layout results differ from the 32-byte path"
        h_hits.zero_()
        r["compact"] = {"value": time_e2e(compact, n1 + n2, reps), "h2d_bytes": 2 * 88, "d2h_bytes": 24 * n1 + n1,
                        "call": "prt_cuda_trace_jobs, PRT_RAYS_PRIMARY + PRT_RAYS_FROM_HITS (camera + seed up; hit records + 1 byte per primary ray down)"}
        dev_hits = Sx["d_hits"].cpu()
        assert torch.equal(h_hits.view(torch.int32), hits_plain.view(torch.int32)) and torch.equal(h_hits.view(torch.int32), dev_hits.view(torch.int32)), \
            "compact-path hit records differ from the 32-byte / device-resident paths"
        hitmask = dev_hits[:, 0] != -1.0
        occ_c = h_occ1[hitmask]
        if mode == PACKET:
            occ_c = occ_c[: n2]
        assert torch.equal(occ_c, occ_plain) and bool((h_occ1[~hitmask] == 2).all()) and torch.equal(occ_plain, Sx["d_occ"].cpu()), "compact-path occlusion results differ"
        if name == "single":                                       # the plain path once more from unpinned (pageable) numpy memory
            pr, ps = h_rays.numpy().copy(), h_sh.numpy().copy()
            ph, po = np.zeros((n1, 6), np.float32), np.zeros(n2, np.uint8)
            pj = make_jobs([dict(kind=0, n=n1, mode=mode, layout=0, rays=pr.ctypes.data, out=ph.ctypes.data), dict(kind=1, n=n2, mode=mode, layout=0, rays=ps.ctypes.data, out=po.ctypes.data)])
            r["pageable"] = {"value": time_e2e(pj, n1 + n2, max(3, reps // 2)), "call": "the plain path from pageable (new[] / numpy) memory"}
            assert np.array_equal(ph.view(np.uint32), hits_plain.numpy().view(np.uint32))
        e2e[name] = r
        del h_rays, h_sh, h_hits, h_occ, h_occ1, h_dirs, h_org

    # ---- secondary rooflines: the tree is cache-resident, so beside the HBM figure the kernel is placed against
    # the measured L2 read bandwidth and the measured FP32 warp-instruction issue rate of this device ----
    l2_peak, issue_peak = ctx.probe_peaks()
    ncu = ncu_constants()

    def roofline(Sx, r, key):
        cn, cp, on, op = TRAVERSAL_CONSTANTS[key]
        b_int = (32.0 * cn + 48.0 * cp + 56.0) * Sx["n1"]
        b_occ = (32.0 * on + 48.0 * op + 33.0) * Sx["n2"]
        dom = "occluded" if r["t_occ"] >= r["t_int"] else "intersect"
        dom_bytes, dom_ms = (b_occ, r["t_occ"]) if dom == "occluded" else (b_int, r["t_int"])
        ach = dom_bytes / (dom_ms * 1e-3) / 1e9
        k = ncu.get(f"{key}_{dom}", {})
        out = {"bound": "issue/latency (scene is cache-resident; the HBM figure below is the contract's denominator, not the limiter)",
               "kernel": f"{dom}KernelDyn<{'packet' if Sx['mode'] == PACKET else 'single'}>", "achieved": ach, "peak": env.hbm_peak, "unit": "GB/s",
               "frac": ach / env.hbm_peak, "traffic": k.get("dram_bytes"), "algorithmic_bytes": dom_bytes,
               "traffic_source": k.get("source"), "peak_source": env.peak_src,
               "l2": {"peak": l2_peak, "unit": "GB/s", "frac": ach / l2_peak, "peak_source": "prt_cuda_probe_peaks: 32 MiB L2-resident buffer, ld.global.cg.v4, all SMs"},
               "kernel_ms": {"intersect": r["t_int"], "occluded": r["t_occ"]},
               "mrays_per_s": {"closest_hit": Sx["n1"] / r["t_int"] / 1e3, "occlusion": Sx["n2"] / r["t_occ"] / 1e3}}
        if k.get("warp_inst"):
            wi = k["warp_inst"] * (Sx["n1"] if dom == "intersect" else Sx["n2"]) / k["rays"]
            out["issue"] = {"achieved": wi / (dom_ms * 1e-3) / 1e9, "peak": issue_peak, "unit": "Gwarp-inst/s", "frac": wi / (dom_ms * 1e-3) / 1e9 / issue_peak,
                            "source": f"smsp__inst_executed.sum per ray ({k.get('source')}) x this launch's rays / live kernel time; peak = prt_cuda_probe_peaks FFMA issue kernel"}
        return out

    rules = {"packet": {"value": rp["value"], "unit": "Mrays/s", "rays_per_pass": P["n1"] + P["n2"], "ms_per_pass": rp["t_int"] + rp["t_occ"],
                        "e2e": e2e["packet"], "roofline": roofline(P, rp, "c2_packet")},
             "single": {"value": rs["value"], "unit": "Mrays/s", "rays_per_pass": S["n1"] + S["n2"], "ms_per_pass": rs["t_int"] + rs["t_occ"],
                        "e2e": e2e["single"], "roofline": roofline(S, rs, "c2_single")},
             "headline": "packet"}

    # ---- what the kernels actually fetch per ray (counting build of the library, one untimed pass) ----
    if rank == 0 and world == 1 and not args.no_fetched and os.path.exists(capi.STATS_LIB_PATH):
        sctx = prt_b200.load_scene(desc, local, stats=True)
        for name, Sx in (("packet", P), ("single", S)):
            sctx.reset_counters()
            sctx.intersect_dev(Sx["n1"], Sx["d_rays"].data_ptr(), Sx["d_avg"].data_ptr() if Sx["d_avg"] is not None else None, Sx["mode"], Sx["d_hits"].data_ptr())
            sctx.occluded_dev(Sx["n2"], Sx["d_sh"].data_ptr(), Sx["mode"], Sx["d_occ"].data_ptr())
            st = sctx.counters()["stats"]
            cls = st["closest_packet" if Sx["mode"] == PACKET else "closest_single"]
            oc = st["occlude"]
            per = lambda c: {"node_records_64B": c["nodeSteps"] / max(c["rays"], 1), "leaf_visits": c["leafVisits"] / max(c["rays"], 1),
                             "triangle_records_48B": c["triTests"] / max(c["rays"], 1),
                             "bytes": (64.0 * c["nodeSteps"] + 48.0 * c["triTests"] + 24.0 * c["uvLoads"]) / max(c["rays"], 1)}
            rules[name]["fetched_per_ray"] = {"closest_hit": per(cls), "occlusion": per(oc),
                                              "source": "libprt_cuda_stats.so (-DPRT_STATS=1) counters, one pass over the same rays"}
        sctx.close()

    # ---- the opt-in fast traversal mode on the same device-resident rays, beside the exact numbers (never the headline) ----
    if rank == 0 and world == 1 and not args.no_fast:
        fctx = prt_b200.load_scene(desc, local, fast=True)
        fctx.set_traversal(True)
        fext = torch.cuda.ExternalStream(fctx.stream(), device=local)
        for name, Sx in (("packet", P), ("single", S)):
            f_hits = torch.empty_like(Sx["d_hits"]); f_occ = torch.empty_like(Sx["d_occ"])
            avg = Sx["d_avg"].data_ptr() if Sx["d_avg"] is not None else None
            reps = 8 if name == "packet" else 32
            tot = 0.0
            fctx.reset_counters()
            with torch.cuda.stream(fext):
                for k in range(reps + 2):
                    env.flush_buf.zero_()
                    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
                    a.record(fext)
                    fctx.intersect_dev(Sx["n1"], Sx["d_rays"].data_ptr(), avg, Sx["mode"], f_hits.data_ptr())
                    fctx.occluded_dev(Sx["n2"], Sx["d_sh"].data_ptr(), Sx["mode"], f_occ.data_ptr())
                    b.record(fext)
                    fctx.synchronize()
                    if k >= 2:
                        tot += a.elapsed_time(b)
            env.launches += int(fctx.counters()["kernelLaunches"])
            fv = (Sx["n1"] + Sx["n2"]) * reps / (tot * 1e-3) / 1e6
            eh, fh = Sx["d_hits"].view(torch.int32), f_hits.view(torch.int32)
            diff = (eh != fh).any(dim=1)
            ties = diff & (eh[:, 0] == fh[:, 0])
            rules[name]["fast_mode"] = {"value": fv, "unit": "Mrays/s", "speedup_over_exact": fv / rules[name]["value"], "passes": reps,
                                        "closest_hit_records_differing": int(diff.sum().item()), "of_which_ties_same_t_other_triangle": int(ties.sum().item()),
                                        "occlusion_results_differing": int((f_occ != Sx["d_occ"]).sum().item()),
                                        "what": "PRT_TRAVERSAL_FAST on the same device-resident rays (prt_cuda_intersect_dev + prt_cuda_occluded_dev)"}
            del f_hits, f_occ
        fctx.close()

    line = {"metric": "Mrays/s (closest-hit + occlusion)", "value": rp["value"], "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": rp["total_ms"] / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": c2_config(S["n1"], S["n2"], P["n1"], P["n2"], passes),
            "timing": {"packet_passes_per_step": max(1, passes // 4), "single_passes_per_step": passes,
                       "timed_region_ms": {"packet": rp["total_ms"], "single": rs["total_ms"]}},
            "e2e": {"value": e2e["packet"]["compact"]["value"], "unit": "Mrays/s", "h2d_bytes_per_step": e2e["packet"]["compact"]["h2d_bytes"],
                    "d2h_bytes_per_step": e2e["packet"]["compact"]["d2h_bytes"], "call": e2e["packet"]["compact"]["call"],
                    "plain": e2e["packet"]["plain"], "shared": e2e["packet"]["shared"],
                    "note": "results of all three forms asserted bit-identical to each other and to the device-resident pass in this run"},
            "rules": rules, "roofline": rules["packet"]["roofline"]}
    line["_ctx"] = (ctx, sets, desc)
    return line


def north_star_primary(env, steps=10):
    """BASELINE.json north_star target: primary-ray closest hit on Cornell+teapot, 1920x1080, 8 spp (one 8-ray packet
    per pixel, Camera::GenerateJitteredRayPacket + Scene::intersect<RayHitPacket,RayPacket> rules).  Rays resident in
    HBM, CUDA events on the context stream, L2 flushed between iterations.  Reported beside the headline, not as it."""
    import prt_b200
    from prt_b200 import scenes
    torch = env.torch
    ctx = prt_b200.load_scene(scenes.cornell_teapot(W2, H2, teapot_refl=scenes.SPECULAR, light=None), env.local)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=env.local)
    n = W2 * H2 * 8
    d_rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    d_avg = torch.empty((n // 8, 3), dtype=torch.float32, device="cuda")
    d_hits = torch.empty((n, 6), dtype=torch.float32, device="cuda")
    ctx.primary_dev((0, 0, W2, H2), 1, 1, prt_b200.PRT_MODE_PACKET, d_rays.data_ptr(), d_avg.data_ptr(), d_hits.data_ptr())
    ctx.synchronize()
    out = {}
    for name, fn in (("intersect_packet", lambda: ctx.intersect_dev(n, d_rays.data_ptr(), d_avg.data_ptr(), prt_b200.PRT_MODE_PACKET, d_hits.data_ptr())),
                     ("intersect_single", lambda: ctx.intersect_dev(n, d_rays.data_ptr(), None, prt_b200.PRT_MODE_SINGLE, d_hits.data_ptr())),
                     ("generate_and_intersect_packet", lambda: ctx.primary_dev((0, 0, W2, H2), 1, 1, prt_b200.PRT_MODE_PACKET, None, None, d_hits.data_ptr()))):
        ms = 0.0
        with torch.cuda.stream(ext):
            for k in range(3 + steps):
                env.flush_buf.zero_()
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record(ext); fn(); e1.record(ext)
                e1.synchronize()
                if k >= 3:
                    ms += e0.elapsed_time(e1)
        out[name] = {"mrays_per_s": n * steps / (ms * 1e-3) / 1e6, "ms": ms / steps}
    hitfrac = float((d_hits[:, 0] != -1.0).float().mean().item())
    ctx.close()
    return {"scene": "Cornell box + teapot 1920x1080, 8 primary rays per pixel", "rays": n, "hit_fraction": hitfrac, **out}


def run_render(env, wl):
    """One render workload through the C ABI: prt_cuda_render_dev (value) / prt_cuda_render (e2e).  At N > 1 both are collective:
    bands per rank, ncclAllGather + de-tile inside the library, the full frame on every rank."""
    import prt_b200
    from prt_b200 import capi
    torch, args, rank, world, local = env.torch, env.args, env.rank, env.world, env.local
    cfg = render_config(wl, args.spp)
    W, H, spp = cfg["W"], cfg["H"], cfg["spp"]
    ctx = prt_b200.load_scene(cfg["desc"], local)
    env.join_comm(ctx)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=local)
    region = (0, 0, W, H)
    fb = torch.zeros((H, W, 3), dtype=torch.float32, device="cuda")
    host_fb = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    cam = cfg["desc"]["camera"]
    steps = args.render_steps or cfg["steps"]
    warm = 1 if (spp > 64 or wl == "c5") else 2

    for k in range(warm):
        ctx.render_dev(region, spp, 100 + k, 1.0, fb.data_ptr())
    ctx.synchronize()
    ctx.reset_counters()
    env.barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(ext):
        e0.record(ext)
        for k in range(steps):
            ctx.render_dev(region, spp, 1 + k, 1.0, fb.data_ptr())
        e1.record(ext)
    env.barrier()
    total_ms = env.reduce(e0.elapsed_time(e1), "MAX")
    cnt = ctx.counters()
    env.launches += int(cnt["kernelLaunches"])
    samples = float(W) * H * (spp // 8 * 8)
    value = samples * steps / (total_ms * 1e-3) / 1e6
    gather_ms = env.reduce(cnt["lastGatherMs"], "MAX") if world > 1 else 0.0

    # the collective alone (no render in front of it, so no waiting for the slowest rank): re-gather the bands of the last frame
    gather_alone_ms = 0.0
    if world > 1:
        reps = 10
        env.barrier()
        g0 = torch.cuda.Event(enable_timing=True); g1 = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            ctx.gather_bands_dev(region, fb.data_ptr())
            g0.record(ext)
            for _ in range(reps):
                ctx.gather_bands_dev(region, fb.data_ptr())
            g1.record(ext)
        env.barrier()
        gather_alone_ms = env.reduce(g0.elapsed_time(g1), "MAX") / reps

    # ---- e2e: prt_cuda_render with a host framebuffer (camera + seed in, RGB32F frame out), seed 1 first: its CRC is the
    # partition-independence check (equal at every N) ----
    e2e_steps = max(1, min(steps, 3))
    L = ctx.L
    import ctypes as C

    def e2e_step(seed):
        ctx.set_camera(cam["pos"], cam["dir"], W, H)             # the per-step host input of a render is camera + seed
        ctx.check(L.prt_cuda_render(ctx.h, 0, 0, W, H, spp, seed, C.c_float(1.0), host_fb.data_ptr()))

    e2e_step(1)
    crc = zlib.crc32(host_fb.numpy().tobytes())
    env.barrier()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        e2e_step(1 + k)
    env.barrier()
    e2e_s = env.reduce(time.perf_counter() - t0, "MAX")
    crcs = [crc]
    if env.dist is not None:
        t = torch.tensor([crc], dtype=torch.int64, device="cuda")
        allc = [torch.zeros_like(t) for _ in range(world)]
        env.dist.all_gather(allc, t)
        crcs = [int(x.item()) for x in allc]

    tot_rays = cnt["primaryRays"] + cnt["extendRays"] + cnt["shadowRays"]
    rec = {"value": value, "unit": "Msamples/s", "ms_per_step": total_ms / steps, "steps": steps, "warmup": warm, "scaling": "strong",
           "config": {"workload": cfg["name"], "spp": spp,
                      "sharding": f"8-row bands dealt round-robin to {world} rank(s), scene replicated, ncclAllGather of the band buffers + de-tile kernel inside libprt_cuda.so" if world > 1
                      else "one GPU: no sharding, no collective",
                      "l2": "per-step working set (wavefront state + queues) exceeds L2"},
           "e2e": {"value": samples * e2e_steps / e2e_s / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": 56, "d2h_bytes_per_step": int(host_fb.numel() * 4),
                   "call": "prt_cuda_set_camera + prt_cuda_render (host RGB32F framebuffer)", "steps": e2e_steps},
           "gather_ms": gather_ms, "gather_alone_ms": gather_alone_ms,
           "gather_note": "gather_ms: all-gather + de-tile of the last timed frame on the slowest rank, including its wait for the other ranks' renders; "
                          "gather_alone_ms: the same collective re-issued on its own" if world > 1 else None,
           "image_crc": f"{crc:08x}", "image_crc_equal_on_all_ranks": len(set(crcs)) == 1,
           "rays_per_step": {"primary": cnt["primaryRays"] * world // steps, "extend": cnt["extendRays"] * world // steps, "shadow": cnt["shadowRays"] * world // steps,
                             "shadow_not_traced_zero_radiance": cnt["shadowRaysSkipped"] * world // steps} if world == 1 else None,
           "mrays_per_s": env.reduce(float(tot_rays), "SUM") / (total_ms * 1e-3) / 1e6}
    k = RENDER_CONSTANTS.get(wl)
    ncu = ncu_constants().get(f"{wl}_wftrace", {})
    if k:
        b = env.reduce((32 * k[0] + 48 * k[1] + 56) * cnt["extendRays"] + (32 * k[2] + 48 * k[3] + 33) * cnt["shadowRays"] + (32 * k[4] + 48 * k[5] + 56) * cnt["primaryRays"], "SUM")
        ach = b / (total_ms * 1e-3) / 1e9 / world
        rec["roofline"] = {"bound": "hbm", "kernel": "wfTrace + wfPrimary (whole step)", "achieved": ach, "peak": env.hbm_peak, "unit": "GB/s", "frac": ach / env.hbm_peak,
                           "traffic": ncu.get("dram_bytes_per_launch"), "traffic_gbs": ncu.get("dram_gbs"),
                           "traffic_per": "the largest wfTrace launch of the ncu capture (a low-spp batch of the same scene); traffic_gbs = that launch's DRAM bytes / its duration",
                           "traffic_source": ncu.get("source"), "peak_source": env.peak_src, "per": "GPU",
                           "l1_data_pipe": ncu.get("l1_data_pipe"), "wftrace_lanes_per_instruction": ncu.get("lanes"),
                           "note": "algorithmic bytes = sum over the step's rays of 32 B per reference node pop + 48 B per triangle + ray / result (per-class constants "
                                   "measured with the oracle's counters) / step time; the trees of c1 / c3 are cache-resident and c4 / c5 are bound by the L1 data pipe "
                                   "(l1_data_pipe, from ncu), so this fraction of the HBM peak is the contract's figure, not a bandwidth-efficiency claim"}
    # what wfTrace fetches per ray class (counting build, a low-spp pass over a band of rows; N = 1 only)
    if rank == 0 and world == 1 and not args.no_fetched and os.path.exists(capi.STATS_LIB_PATH) and wl != "c5":
        sctx = prt_b200.load_scene(cfg["desc"], local, stats=True)
        y0 = H // 4
        tmp = torch.zeros((H, W, 3), dtype=torch.float32, device="cuda")
        sctx.reset_counters()
        sctx.render_dev((0, y0, W, min(H, y0 + max(8, H // 8))), 8, 1, 1.0, tmp.data_ptr())
        st = sctx.counters()["stats"]
        rec["fetched_per_ray"] = {c: {"node_records_64B": v["nodeSteps"] / max(v["rays"], 1), "triangle_records_48B": v["triTests"] / max(v["rays"], 1),
                                      "bytes": (64.0 * v["nodeSteps"] + 48.0 * v["triTests"] + 24.0 * v["uvLoads"]) / max(v["rays"], 1), "rays": v["rays"]}
                                  for c, v in st.items() if v["rays"]}
        rec["fetched_per_ray"]["source"] = "libprt_cuda_stats.so counters of wfTrace (extension + shadow rays; the primary pass is not counted), 8 spp on an eighth of the rows"
        sctx.close()
        if not args.no_fast:                                       # the same band in the fast traversal mode: 96-byte wide-node records, 48-byte triangles
            fsctx = prt_b200.load_scene(cfg["desc"], local, stats=True, fast=True)
            fsctx.set_traversal(True)
            fsctx.reset_counters()
            fsctx.render_dev((0, y0, W, min(H, y0 + max(8, H // 8))), 8, 1, 1.0, tmp.data_ptr())
            fst = fsctx.counters()["stats"]
            rec["fetched_per_ray"]["fast_mode"] = {c: {"wide_node_records_96B": v["nodeSteps"] / max(v["rays"], 1), "triangle_records_48B": v["triTests"] / max(v["rays"], 1),
                                                       "bytes": (96.0 * v["nodeSteps"] + 48.0 * v["triTests"] + 24.0 * v["uvLoads"]) / max(v["rays"], 1), "rays": v["rays"]}
                                                   for c, v in fst.items() if v["rays"]}
            fsctx.close()
        del tmp
    # the opt-in fast traversal mode (PRT_OPTION_TRAVERSAL = PRT_TRAVERSAL_FAST, prt_wide.cuh) beside the exact one: same call,
    # same frame; not the headline (it does not promise the reference's tie-breaks), so it never replaces `value`
    if rank == 0 and world == 1 and not args.no_fast and wl != "c1":
        exact_img = host_fb.numpy().copy() if e2e_steps == 1 else None
        if exact_img is None:
            e2e_step(1); exact_img = host_fb.numpy().copy()
        ctx.close()
        fctx = prt_b200.load_scene(cfg["desc"], local, fast=True)
        fctx.set_traversal(True)
        fext = torch.cuda.ExternalStream(fctx.stream(), device=local)
        fsteps = max(2, steps // 2)
        fctx.render_dev(region, spp, 100, 1.0, fb.data_ptr()); fctx.synchronize()
        f0 = torch.cuda.Event(enable_timing=True); f1 = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(fext):
            f0.record(fext)
            for k in range(fsteps):
                fctx.render_dev(region, spp, 1 + k, 1.0, fb.data_ptr())
            f1.record(fext)
        fctx.synchronize()
        fms = f0.elapsed_time(f1)
        env.launches += int(fctx.counters()["kernelLaunches"])
        fctx.render_dev(region, spp, 1, 1.0, fb.data_ptr()); fctx.synchronize()
        fast_img = fb.cpu().numpy()
        fv = samples * fsteps / (fms * 1e-3) / 1e6
        rec["fast_mode"] = {"value": fv, "unit": "Msamples/s", "steps": fsteps, "ms_per_step": fms / fsteps, "speedup_over_exact": fv / value,
                            "pixels_differing_from_exact": float((fast_img != exact_img).any(axis=2).mean()),
                            "relative_l1_difference": float(np.abs(fast_img.astype(np.float64) - exact_img).sum() / max(float(np.abs(exact_img).sum()), 1e-9)),
                            "what": "prt_cuda_set_option(PRT_OPTION_TRAVERSAL, PRT_TRAVERSAL_FAST): 8-wide nodes with 8-bit quantised boxes over a tree with 4-triangle "
                                    "leaves; same closest t and triangle arithmetic, ties between coplanar triangles and grazing rejections not promised "
                                    "(tests/test_gpu_fast_mode.py enumerates them)"}
        fctx.close()
    else:
        ctx.close()
    del fb
    torch.cuda.empty_cache()
    return rec, cfg


def cpu_baseline_c2(line):
    """Reference (or oracle port) timed on this box's host cores on a bounded sample of the same workload, and the ID-agreement
    check of the headline rays: our device-resident hit records against the reference's, ray by ray."""
    ref = RefC2()
    ctx, sets, desc = line["_ctx"]
    n1s, n2s, n1p, n2p = ref.counts()
    vp, reps_p, dt_p = ref.timed(ref.pass_packet, n1p + n2p, budget_s=6.0)
    vs, reps_s, dt_s = ref.timed(ref.pass_single, n1s + n2s, budget_s=6.0)
    parity = {}
    for name, rhits in (("packet", ref.p_hits), ("single", ref.s_hits)):
        g = sets[name]["d_hits"].cpu().numpy().view(np.uint32)
        r = rhits.view(np.uint32)
        hitmiss = int(((g[:, 0] == 0xbf800000) != (r[:, 0] == 0xbf800000)).sum())
        ids = int(((g[:, 4:6] != r[:, 4:6]).any(axis=1)).sum())
        bits = int((g != r).any(axis=1).sum())
        parity[name] = {"rays": int(r.shape[0]), "hit_miss_mismatches": hitmiss, "id_mismatches": ids, "record_bit_mismatches": bits}
    line["parity"] = {"against": f"{ref.kind} (seed 1 frame, closest-hit records compared ray by ray)", **parity}
    return {"value": vp, "unit": "Mrays/s", "cores": ref.threads, "kind": ref.kind,
            "sample": f"{reps_p} passes over the frame's {n1p + n2p} packet-rule rays in {dt_p:.1f} s",
            "single_rules": {"value": vs, "sample": f"{reps_s} passes over {n1s + n2s} single-rule rays in {dt_s:.1f} s"}}


def run_ours(args):
    env = Env(args)
    rank, world, local = env.rank, env.world, env.local
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    line = {}
    if args.workload in ("all", "c2"):
        line = run_c2(env, sampler)
    if args.workload == "all":
        wls = ["c1", "c3", "c4", "c5"] if world == 1 else ["c3", "c5"]     # N > 1: the configs BASELINE.json quotes at 1/2/4/8 GPUs
    else:
        wls = [] if args.workload == "c2" else [args.workload]
    render = {}
    from oracle import refprt   # cpu_baseline legs only
    for wl in wls:
        rec, cfg = run_render(env, wl)
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            kind = "reference" if refprt.available() else "port"
            v, what = ref_render_sample(cfg, threads, kind, refprt, repeats=3 if wl == "c1" else 1)
            rec["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": threads, "kind": kind, "sample": what}
        render[wl] = rec
    clocks = sampler.stop() if rank == 0 else None
    if "metric" not in line:                      # a single render workload was asked for: it is the line's metric
        wl = wls[0]
        rec = render[wl]
        line = {"metric": "Msamples/s (full path tracing)", "value": rec["value"], "unit": "Msamples/s", "n_gpus": world, "steps": rec["steps"], "warmup": rec["warmup"],
                "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": rec["config"], "e2e": rec["e2e"], "roofline": rec.get("roofline"), "cpu_baseline": rec.get("cpu_baseline")}
    if render:
        line["render"] = render
    if "_ctx" in line:
        if rank == 0 and world == 1:
            line["north_star_primary"] = north_star_primary(env)
        if rank == 0 and not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline_c2(line)
        line.pop("_ctx")[0].close()
    line["gpu_launches"] = int(env.launches)
    line["clocks"] = clocks
    if rank == 0:
        print(json.dumps(line), flush=True)
    if env.dist is not None:
        env.dist.barrier()
        env.dist.destroy_process_group()


def main():
    args = parse()
    # stdout carries exactly one JSON line: anything a library prints there (NCCL's version banner, the reference's build log)
    # goes to stderr instead
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.impl == "reference":
        run_reference(args, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return
    run_ours(args)


if __name__ == "__main__":
    main()
