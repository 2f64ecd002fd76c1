#!/usr/bin/env python
"""bench.py — one JSON line per run (contract in the task statement, tier reading).

A step = one pass of the hot path (raygen -> [extend -> shade] x depth -> accumulate) over one frame of the
headline workload, BASELINE.json configs[1]: procedural 262 144-triangle atrium, 1024x1024, 64 spp, max depth 5,
one B200.  With N GPUs every rank renders its own 64-sample range of the same frame (disjoint Philox sample
indices, replicated BVH) and the f32 accumulators are summed on rank 0 with one NCCL reduce: weak scaling.

  value      counted Mrays/s (one ray = one extend invocation) of the whole job, scene + BVH resident in HBM
  e2e        the same through the C-ABI with HOST buffers: cupt_set_scene + cupt_set_bvh (H2D of the scene arrays
             and the BVH image), cupt_render, cupt_read_accum (D2H of the image) inside the timed region
  roofline   extend kernel: ALGORITHMIC BVH bytes (node visits x 80 B + triangle tests x 48 B, SURVEY 8d, counted by a
             stats pass of the same step) / its CUDA-event time, against the measured HBM copy bandwidth; the 44 B per
             ray of ray / hit / queue records and ncu's DRAM bytes are separate fields.  `bound` says what ncu shows
             limits the kernel (instruction issue: the 16 MB BVH is L1/L2-resident), not what the fraction is taken of
  strong     (every line) BASELINE config 3's 3840x2160 frame with a FIXED total of 128 spp split over the ranks by sample
             ranges: device ms and e2e ms, so that the per-N lines give a strong-scaling curve next to the weak one
  multi_gpu_check  (N > 1) rank 0 re-renders every other rank's sample range for a band of image tiles on top of its own
             and compares with the NCCL-summed image
  cpu_baseline  the reference's own integrator (oracle/_ref, else the oracle port) on the host cores, bounded sample
  --impl reference   times that CPU implementation as the reference arm
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

W, H, SPP, DEPTH, SEED, NTRIS = 1024, 1024, 64, 5, 1234, 262144
WORKLOAD = f"atrium {NTRIS} tris (procedural, seed 1), {W}x{H}, {SPP} spp, max depth {DEPTH}, seed {SEED}"
METRIC = "Mrays/s (counted rays; Sponza-scale scene 1024x1024)"


def build_scene():
    from cudapathtracer_b200 import api, scenes
    sc = scenes.atrium(NTRIS)
    s = api.SceneHandle()
    s.add_scene(sc)
    cam = api.make_camera(W, H, **sc.camera)
    return sc, s, cam


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.t0, self.t1 = index, [], None, None, None

    def mark_begin(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "25",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            if self.t0 is not None and self.t1 is None:   # keep only what was sampled inside the timed region
                self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_summary():
    """the committed ncu --set full summary of the extend kernel (profiles/extend_ncu_summary.json, written by
    tools/ncu_summarise.py from one capture of this same command): per-launch DRAM bytes and what limits the kernel"""
    p = os.path.join(ROOT, "profiles", "extend_ncu_summary.json")
    try:
        d = json.load(open(p))
    except Exception:
        return {}
    la = d.get("launches") or []
    w = [x.get("ms", 0.0) for x in la]
    def avg(key):
        v = [(x[key], m) for x, m in zip(la, w) if x.get(key) is not None]
        return sum(a * m for a, m in v) / sum(m for _, m in v) if v and sum(m for _, m in v) > 0 else None
    return {"dram_bytes_per_launch": d.get("dram_bytes_per_launch"), "generation": d.get("generation"),
            "threads_per_inst": avg("threads_per_inst"), "issue_active_pct": avg("issue_active_pct"),
            "warps_active_pct": avg("warps_active_pct"), "l1_hit_pct": avg("l1_hit_pct"), "l2_hit_pct": avg("l2_hit_pct"),
            "ms_per_launch_under_ncu": (sum(w) / len(w)) if w else None}


class _Quiet:
    """the reference printf()s while it loads and builds (modelLoader.h:134-209, BVH.h:379-380,445,464)"""

    def __enter__(self):
        sys.stdout.flush()
        self.devnull, self.saved = os.open(os.devnull, os.O_WRONLY), os.dup(1)
        os.dup2(self.devnull, 1)

    def __exit__(self, *a):
        os.dup2(self.saved, 1); os.close(self.devnull); os.close(self.saved)


_ref_loaded = [None]


def cpu_reference(cam, spp, threads):
    """the reference's integrator on host cores: (rays, seconds, kind, BVH build seconds).

    oracle/_ref (the real reference compiled for the host) when it was built: the scene goes in through the reference's
    OWN loadOBJ (modelLoader.h:125) from the OBJ/MTL text of the procedural atrium and its own buildBVH() — nothing of
    libcupt.so is on this path.  Only without oracle/_ref: the oracle's C port, which needs the reference-layout BVH2
    and therefore the product's host builder of it (cupt_bvh2_build, pinned bit for bit in tests/test_bvh2.py)."""
    import refbind
    from cudapathtracer_b200 import scenes
    if refbind.available():
        if _ref_loaded[0] is None:
            obj, mtl = scenes.to_obj(scenes.atrium(NTRIS))
            with _Quiet():
                refbind.reset()
                refbind.load_obj_text(obj, mtl)
                t0 = time.perf_counter(); refbind.build_bvh(); _ref_loaded[0] = time.perf_counter() - t0
        t0 = time.perf_counter()
        _, _, rays = refbind.render(cam, integrator=2, num_bounces=DEPTH, rng=1, seed=SEED, first_sample=0, num_samples=spp, nthreads=threads)
        return rays, time.perf_counter() - t0, "reference", _ref_loaded[0]
    import oraclebind
    _, sh, _ = build_scene()
    v, t, m, l = sh.arrays()
    n2, d2 = sh.build_bvh2()
    osc = oraclebind.OracleScene(v, t, m, l, sh.total_light_area, n2)
    t0 = time.perf_counter()
    _, _, rays = osc.render_pt(cam, DEPTH, SEED, 0, spp)
    return rays, time.perf_counter() - t0, "port", 0.0


def make_camera_np(width, height):
    """the reference's camera POD (camera.h:26-35) without the product library: same fields as api.make_camera"""
    from cudapathtracer_b200 import api, scenes   # (importing the binding module does not load libcupt.so)
    return api.make_camera(width, height, **scenes.atrium_camera())


def legacy_gpu(scene_handle, cam, spp=16):
    """reported baseline: the reference's own kernel.cu rebuilt for sm_100a (oracle/_ref/libcupt_legacy.so), its live
    integrator (fixed-length BDPT, kernel.cu:217-415), 512-thread blocks, one launch + device sync per sample"""
    try:
        import legacybind
        if not legacybind.available() or legacybind.image_dim() != W:
            return {"unavailable": "oracle/_ref/libcupt_legacy.so not built for this image size"}
        v, t, m, l = scene_handle.arrays()
        r = legacybind.render(v, t, m, l, scene_handle.total_light_area, cam, spp)
        return {"value": r["rays"] / (r["ms"] * 1e-3) / 1e6, "unit": "Mrays/s", "msamples_per_s": W * H * spp / (r["ms"] * 1e-3) / 1e6,
                "ms": r["ms"], "spp": spp, "rays_per_sample": r["rays"] / (W * H * spp), "bvh_depth": r["depth"], "bvh_build_s": r["build_s"],
                "what": "original kernel.cu, compile patches only, -arch=sm_100a, drawPixel (BDPT) per sample with cudaDeviceSynchronize; rays counted in a second, untimed run of the same length"}
    except Exception as e:   # a reported baseline must not take the bench line down
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cam = make_camera_np(W, H)
    threads = os.cpu_count() or 1
    spp = 1   # bounded sample per step: 1024x1024 x 1 spp of the same frame
    times, rays_total = [], 0
    kind = "port"
    for i in range(args.warmup + args.steps):
        rays, dt, kind, _ = cpu_reference(cam, spp, threads)
        if i >= args.warmup:
            times.append(dt); rays_total += rays
    total = sum(times)
    val = rays_total / total / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64/f32 (reference: f32 geometry, f64 colour)", "data": "synthetic",
            "config": {"workload": WORKLOAD},
            "step": f"{W}x{H} x {spp} spp of the frame on host cores, scene through the reference's own loadOBJ + buildBVH",
            "msamples_per_s": W * H * spp * args.steps / total / 1e6,
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": kind,
                             "sample": f"{W}x{H} x {spp} spp per step, radianceAlongSingleStep2 depth {DEPTH}, keyed Philox"},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0

STRONG_W, STRONG_H, STRONG_SPP = 3840, 2160, 128


def multi_gpu_check(ctx, api, sharding, dist, torch, rank, world, ranges, accum_t, cam_w, cam_h):
    """Does the NCCL-summed image of ranks rendering their own sample ranges equal ONE GPU rendering all of them?
    Every rank renders its range and the accumulators are reduced to rank 0 (exactly the timed step); rank 0 then
    renders every OTHER rank's range on top of its own accumulator for a band of image tiles (cupt_set_tile_window)
    and compares that band with the reduced image.  Philox is keyed by the global sample index, so the two differ
    only by the f32 summation order: required <= 1e-6 relative."""
    first, n = ranges[rank]
    ctx.set_tile_window(0, 0)
    ctx.reset_accum(); ctx.render(first, n)
    ctx.copy_accum_device(accum_t.data_ptr())
    torch.cuda.synchronize()
    dist.reduce(accum_t, dst=0, op=dist.ReduceOp.SUM)
    torch.cuda.synchronize()
    out = None
    if rank == 0:
        reduced = accum_t.view(cam_h, cam_w, 3).cpu().numpy()
        ntiles = ctx.tile_count()
        tiles_x = (cam_w + 7) // 8
        rows = max(1, (cam_h // 4) // 16)                       # a band of 1/16 of the tile rows, in the middle of the image
        row0 = (cam_h // 4) // 2 - rows // 2
        ctx.set_tile_window(row0 * tiles_x, rows * tiles_x)
        for r in range(1, world):
            ctx.render(*ranges[r])
        ctx.set_tile_window(0, 0)
        own = ctx.read_accum(False).reshape(cam_h, cam_w, 3)
        y0, y1 = row0 * 4, (row0 + rows) * 4
        a, b = own[y0:y1].astype(np.float64), reduced[y0:y1].astype(np.float64)
        denom = np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-3)
        rel = np.abs(a - b) / denom
        out = {"max_abs_diff": float(np.abs(a - b).max()), "max_rel_diff": float(rel.max()), "pixels": int((y1 - y0) * cam_w),
               "band_rows": [int(y0), int(y1)], "mean_radiance_sum": float(b.mean()), "ranks_rerendered": world - 1,
               "ok": bool(rel.max() <= 1e-6 and b.mean() > 0),
               "what": "rank 0: own accumulator + every other rank's sample range re-rendered for the band, vs the NCCL-reduced image"}
        assert ntiles == tiles_x * ((cam_h + 3) // 4)
    dist.barrier()
    return out


def strong_record(ctx, api, sharding, dist, torch, s, bvh, sc, rank, world, stream, steps, group=None):
    """BASELINE config 3's frame (3840x2160) with a FIXED total of STRONG_SPP samples split over the ranks by sample
    ranges: device ms (CUDA events, max over ranks) and e2e ms (host buffers: scene + BVH H2D on every rank, render,
    NCCL reduce, image D2H on rank 0)."""
    cam = api.make_camera(STRONG_W, STRONG_H, **sc.camera)
    ctx.set_camera(cam)
    ranges = [sharding.sample_range(r, world, STRONG_SPP) for r in range(world)]
    first, n = ranges[rank]
    accum = torch.zeros(STRONG_W * STRONG_H * 3, dtype=torch.float32, device="cuda")
    host_t = torch.empty((STRONG_W * STRONG_H, 3), dtype=torch.float32, pin_memory=True)

    def step():
        ctx.reset_accum(); ctx.render(first, n)
        if world > 1:
            ctx.copy_accum_device(accum.data_ptr())
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)

    def e2e_step():
        if group is not None:   # the public multi-GPU API: scene + BVH staged on rank 0 and broadcast, one ncclReduce, image read on rank 0
            group.set_scene(s if rank == 0 else None); group.set_camera(cam); group.set_bvh(bvh if rank == 0 else None)
            group.reset_accum(); group.render(0, STRONG_SPP)
            if rank == 0:
                group.read_accum(False, out=host_t.numpy())
            else:
                group.sync()
            return
        ctx.set_scene(s); ctx.set_camera(cam); ctx.set_bvh(bvh)
        ctx.reset_accum(); ctx.render(first, n)
        if world > 1:
            ctx.copy_accum_device(accum.data_ptr())
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                host_t.copy_(accum.view(-1, 3))
        else:
            ctx.read_accum(False, out=host_t.numpy())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx.reset_stats(); step(); barrier()              # warm-up (allocates the 4K path buffers) + ray count of one step
    rays = torch.tensor([float(ctx.stats()["rays"])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    step(); barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        a.record(stream); step(); b.record(stream)
    barrier()
    tms = torch.tensor([sum(a.elapsed_time(b) for a, b in ev) / steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    e2e_step(); barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    barrier()
    te = torch.tensor([(time.perf_counter() - t0) / steps * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    check = multi_gpu_check(ctx, api, sharding, dist, torch, rank, world, ranges, accum, STRONG_W, STRONG_H) if world > 1 else None
    ms, e2e_ms, r = float(tms.item()), float(te.item()), float(rays.item())
    rec = {"workload": f"atrium {NTRIS} tris, {STRONG_W}x{STRONG_H}, {STRONG_SPP} spp in total split over {world} GPU(s) by sample ranges, max depth {DEPTH}",
           "scaling": "strong", "ms": ms, "e2e_ms": e2e_ms, "steps": steps, "rays_per_step": r,
           "value": r / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "e2e_value": r / (e2e_ms * 1e-3) / 1e6,
           "msamples_per_s": STRONG_W * STRONG_H * STRONG_SPP / (ms * 1e-3) / 1e6,
           "samples_per_rank": [int(x[1]) for x in ranges], "reduce_bytes": STRONG_W * STRONG_H * 12 if world > 1 else 0}
    if check is not None:
        rec["multi_gpu_check"] = check
    del accum, host_t
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the `strong` sub-record (config 3's 4K frame, 128 spp split over the GPUs)")
    ap.add_argument("--traversal", default="wide", choices=["wide", "strict", "ref"], help="wide (default), strict = wide + reference-equality check, ref = the reference algorithm on its own BVH2")
    ap.add_argument("--integrator", default="pt", choices=["pt", "bdpt"], help="pt = radianceAlongSingleStep2 with max depth 5 (the metric's config); bdpt = the reference's live radianceAlongSingleStep (what the legacy kernel.cu runs)")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3], help="BASELINE.json config: 2 (default: 1024^2, 64 spp per GPU, weak scaling) or 3 (3840x2160, 1024 spp split over the GPUs, strong scaling)")
    ap.add_argument("--tri-test", default="refexact", choices=["refexact", "watertight"], help="refexact = the reference's triIntersect bit for bit (default); watertight = Woop et al. on exact vertices")
    ap.add_argument("--share-primary", action="store_true", help="trace the (sample-independent) pinhole primary ray once per pixel and pass; fewer rays are traced AND counted, so Mrays/s drops while Msamples/s rises")
    ap.add_argument("--partition", default="samples", choices=["samples", "tiles"], help="multi-GPU split of a config 3 frame: sample ranges (default) or image tiles (8x4-pixel tiles, contiguous range per rank, every rank renders all samples of its tiles); both end in the same NCCL sum of the f32 accumulators")
    ap.add_argument("--bvh", default="host", choices=["host", "device"], help="host = binned-SAH builder (OpenMP), uploaded; device = cupt_build_bvh_device, rebuilt inside every e2e step")
    args = ap.parse_args()
    global W, H, SPP, WORKLOAD
    if args.config == 3:
        W, H, SPP = 3840, 2160, 1024
        WORKLOAD = f"atrium {NTRIS} tris (procedural, seed 1), {W}x{H}, {SPP} spp over all GPUs, max depth {DEPTH}, seed {SEED}"
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    from cudapathtracer_b200 import api

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    sc, s, cam = build_scene()
    mode = {"wide": api.TRAVERSAL_WIDE, "strict": api.TRAVERSAL_WIDE_STRICT, "ref": api.TRAVERSAL_REF_BVH2}[args.traversal]
    # reference-layout tree: the tie-break ranks of every mode (equal-t hits resolve like the reference's left-first
    # visiting order) and the structure the strict / reference-exact modes check against
    n2, d2 = s.build_bvh2()
    t0 = time.perf_counter()
    bvh = s.build_bvh(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]))
    build_s = time.perf_counter() - t0
    info = bvh.info()
    ctx = api.Context(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    ctx.set_scene(s); ctx.set_bvh(bvh); ctx.set_camera(cam)
    dev_rank = api.bvh2_leaf_order(n2, s.counts()[1])
    dev_info = None
    if args.bvh == "device":
        ctx.build_bvh_device(tie_rank=dev_rank)
        dev_info = ctx.build_bvh_device(tie_rank=dev_rank)
        info = dict(info, num_nodes=dev_info["num_nodes"], depth=dev_info["depth"])
    if mode != api.TRAVERSAL_WIDE:
        ctx.set_bvh2(n2, d2)
    ctx.set_options(max_depth=DEPTH, seed=SEED, traversal=mode, time_stages=1, collect_stats=0,
                    integrator=api.INTEGRATOR_BDPT if args.integrator == "bdpt" else api.INTEGRATOR_PT,
                    tri_test=api.TRI_WATERTIGHT if args.tri_test == "watertight" else api.TRI_REFEXACT,
                    share_primary_hits=int(args.share_primary))
    accum_t = torch.zeros(W * H * 3, dtype=torch.float32, device="cuda")
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    from cudapathtracer_b200 import sharding
    if args.partition == "tiles" and args.config != 3:
        raise SystemExit("--partition tiles splits ONE frame over the ranks: use it with --config 3 (strong scaling)")
    if args.config == 3 and args.partition == "tiles":   # strong scaling by image tiles: disjoint pixels, all samples each
        first, spp_rank = 0, SPP
        ctx.set_tile_window(*sharding.tile_range(rank, world, ctx.tile_count()))
        samples_step = W * H * SPP
    elif args.config == 3:   # strong scaling: the frame's samples are split over the ranks
        first, spp_rank = sharding.sample_range(rank, world, SPP)
        samples_step = W * H * SPP
    else:
        first, _ = sharding.weak_sample_range(rank, SPP)
        spp_rank = SPP
        samples_step = W * H * SPP * world

    def step():
        ctx.reset_accum()
        ctx.render(first, spp_rank)
        if world > 1:
            ctx.copy_accum_device(accum_t.data_ptr())
            sharding.reduce_accumulators(accum_t, dst=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # stats pass (untimed): node visits / triangle tests / rays of exactly this step
    ctx.set_options(collect_stats=1)
    ctx.reset_stats(); step(); torch.cuda.synchronize()
    st = ctx.stats()
    rays_step, nodes_step, tris_step = st["rays"], st["nodes_visited"], st["tris_tested"]
    ctx.set_options(collect_stats=0)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()          # nvidia-smi takes a moment to come up: started before the warm-up, rows kept only inside the timed region
    for _ in range(args.warmup):
        step()
    barrier()
    ctx.reset_stats()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    sampler.mark_begin()
    for a, b in ev:
        flush_buf.fill_(1)          # L2 flush between timed iterations (outside the event pair)
        a.record(stream); step(); b.record(stream)
    barrier()
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None
    ms = sum(a.elapsed_time(b) for a, b in ev)
    st = ctx.stats()
    tms = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_total = float(tms.item())
    # every rank traces a different sample range of the same frame: sum the rays each rank actually counted
    trays = torch.tensor([float(st["rays"])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(trays, op=dist.ReduceOp.SUM)
    rays_all = float(trays.item())
    value = rays_all / (ms_total * 1e-3) / 1e6

    # e2e through the C-ABI with host buffers (scene + BVH H2D, render, image D2H), same step
    host_t = torch.empty((W * H, 3), dtype=torch.float32, pin_memory=True)   # the caller's image buffer: pinned, so the read-back is one DMA
    host_img = host_t.numpy()
    nv, nt, nm, nl, _ = s.counts()
    h2d = nv * 12 + nt * 28 + nt * 16 + nm * 56 + (nl + 1) * 48 + 32
    if args.bvh == "host":
        h2d += info["num_nodes"] * 80 + info["num_tri_records"] * 48      # the device builder uploads no tree
    if mode != api.TRAVERSAL_WIDE:
        h2d += len(n2) * 32 * 2 + nt * 4                                   # reference-layout tree, its parent-linked copy, leaf parents
    d2h = W * H * 12

    # N > 1: the public multi-GPU API, cupt_group (one member per process here, torchrun's shape): the scene and the BVH are staged
    # through host memory on rank 0 only and reach the other ranks by ncclBroadcast; one ncclReduce; the image is read on rank 0
    group = None
    if world > 1 and args.partition == "samples":
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(api.group_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, src=0)
        group = api.Group(device=local, rank=rank, nranks=world, unique_id=bytes(uid.cpu().numpy().tobytes()))
        group.set_options(max_depth=DEPTH, seed=SEED, traversal=mode, integrator=api.INTEGRATOR_BDPT if args.integrator == "bdpt" else api.INTEGRATOR_PT,
                          tri_test=api.TRI_WATERTIGHT if args.tri_test == "watertight" else api.TRI_REFEXACT, share_primary_hits=int(args.share_primary))
    e2e_total_spp = SPP if args.config == 3 else SPP * world

    def e2e_step():
        if group is not None:
            root = rank == 0
            group.set_scene(s if root else None); group.set_camera(cam)
            if args.bvh == "device":
                group.build_bvh_device(tie_rank=dev_rank)
            else:
                group.set_bvh(bvh if root else None)
            if mode != api.TRAVERSAL_WIDE:
                group.set_bvh2(n2 if root else None, d2)
            group.reset_accum(); group.render(0, e2e_total_spp)
            if root:
                group.read_accum(False, out=host_img)
            else:
                group.sync()
            return
        ctx.set_scene(s); ctx.set_camera(cam)
        if args.bvh == "device":
            ctx.build_bvh_device(tie_rank=dev_rank)      # the build itself is inside the e2e step
        else:
            ctx.set_bvh(bvh)
        if mode != api.TRAVERSAL_WIDE:
            ctx.set_bvh2(n2, d2)
        ctx.reset_accum(); ctx.render(first, spp_rank)
        if world > 1:
            ctx.copy_accum_device(accum_t.data_ptr())
            dist.reduce(accum_t, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                host_t.copy_(accum_t.view(-1, 3))
        else:
            ctx.read_accum(False, out=host_img)

    e2e_step(); barrier()
    t0 = time.perf_counter()
    n_e2e = max(2, min(args.steps, 5))
    for _ in range(n_e2e):
        e2e_step()
    barrier()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = rays_step * world * n_e2e / float(te.item()) / 1e6

    # collective extras (every rank takes part): correctness of the N-GPU sum, and the strong-scaling record
    mgc = None
    plain = mode == api.TRAVERSAL_WIDE and args.integrator == "pt" and args.partition == "samples"
    if world > 1 and plain:
        ranges = [(sharding.weak_sample_range(r, SPP) if args.config == 2 else sharding.sample_range(r, world, SPP)) for r in range(world)]
        mgc = multi_gpu_check(ctx, api, sharding, dist, torch, rank, world, ranges, accum_t, W, H)
    strong = None
    if args.config == 2 and plain and not args.no_strong and args.bvh == "host":
        strong = strong_record(ctx, api, sharding, dist, torch, s, bvh, sc, rank, world, stream, max(1, min(args.steps, 2)), group)
        ctx.set_camera(cam)

    if rank == 0:
        peak, peak_src = measured_peak()
        n_ext = max(1, st["extend_launches"])
        ext_ms = st["ms_extend"]
        # ALGORITHMIC bytes per ray = node visits x 80 B + primitive tests x 48 B (the layout shipped, csrc/layout.h);
        # the wavefront's ray / hit / queue records (32 B ray read + 8 B hit write + 4 B queue entry) are a separate term
        bytes_step = nodes_step * 80 + tris_step * 48
        queue_bytes_per_ray = 44
        achieved = bytes_step * args.steps / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None
        ncu = ncu_summary()
        traffic = ncu.get("dram_bytes_per_launch")
        launch_ms = ext_ms / n_ext
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong" if args.config == 3 else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "traversal": args.traversal, "integrator": args.integrator, "tri_test": args.tri_test, "share_primary_hits": bool(args.share_primary), "parallelism": f"{'image tiles' if args.partition == 'tiles' else 'sample ranges'} x{world}, replicated BVH, NCCL reduce of f32 accumulators",
                       "l2": "L2 flushed (256 MiB write) between timed steps; per-step path state (64 Mi paths x 72 B = 4.8 GB) exceeds the 126 MB L2, the 16 MB BVH is L2-resident within a step by design",
                       "bvh": {"builder": args.bvh, "nodes": info["num_nodes"], "prims": info["num_tri_records"], "depth": info["depth"], "build_s_host": build_s,
                               "build_ms_device": None if dev_info is None else {k: v for k, v in dev_info.items() if k.startswith("ms_")}}},
            "msamples_per_s": samples_step * args.steps / (ms_total * 1e-3) / 1e6,
            "rays_per_step_per_gpu": rays_step, "rays_per_sample": rays_step / (W * H * spp_rank),
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n_e2e,
                    "api": "cupt_group (rank 0 stages scene + BVH once, ncclBroadcast to the other ranks, ncclReduce, image D2H on rank 0)" if group is not None
                           else ("cupt_context per rank + torch.distributed reduce" if world > 1 else "cupt_context")},
            "gpu_launches": int(st["kernel_launches"]),
            "stage_ms_per_step": {k: st["ms_" + k] / args.steps for k in ("raygen", "extend", "shade", "accum")},
            "roofline": {"bound": "issue", "bound_note": "instruction issue / L1-L2 latency: the 16 MB BVH is cache-resident, DRAM carries the path records only "
                                  "(ncu fields below); `achieved` / `frac` are SURVEY 8d's ALGORITHMIC BVH-fetch bytes against the HBM copy peak, not DRAM traffic",
                         "roofline_of": "hbm", "kernel": "k_extend<0,false,*>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": (achieved / peak) if achieved else None, "peak_source": peak_src, "traffic": traffic,
                         "dram_frac": (traffic / (launch_ms * 1e-3) / 1e9 / peak) if traffic and launch_ms > 0 else None,
                         "algorithmic_bytes_per_launch": bytes_step / (n_ext / args.steps) if n_ext else None,
                         "launch_ms": launch_ms, "launches": int(n_ext),
                         "bytes_per_ray": bytes_step / rays_step, "nodes_per_ray": nodes_step / rays_step, "tris_per_ray": tris_step / rays_step,
                         "node_bytes": 80, "prim_bytes": 48, "queue_bytes_per_ray": queue_bytes_per_ray,
                         "ncu": {k: ncu.get(k) for k in ("generation", "issue_active_pct", "threads_per_inst", "warps_active_pct", "l1_hit_pct", "l2_hit_pct", "ms_per_launch_under_ncu")}},
        }
        if mgc is not None:
            line["multi_gpu_check"] = mgc
        if strong is not None:
            line["strong"] = strong
        if world == 1 and mode == api.TRAVERSAL_WIDE and not args.share_primary and args.tri_test == "refexact":
            # the same step in the mode whose hits equal the reference's trace() on EVERY ray (the plain wide mode
            # returns the nearest triangle also on the ~1 ray in 40 000 where the reference's slab test loses it)
            ctx.set_bvh2(n2, d2)
            ctx.set_options(traversal=api.TRAVERSAL_WIDE_STRICT)
            step(); torch.cuda.synchronize()
            ctx.reset_stats()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flush_buf.fill_(1); e0.record(stream)
            for _ in range(args.steps):
                step()
            e1.record(stream); torch.cuda.synchronize()
            sst = ctx.stats()
            ms_s = e0.elapsed_time(e1)
            line["strict_mode"] = {"value": sst["rays"] / (ms_s * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms_s / args.steps,
                                   "retraced_rays_per_step": sst["strict_retraces"] / args.steps,
                                   "what": "CUPT_TRAVERSAL_WIDE_STRICT: hits bit-identical to the reference's on every ray"}
            ctx.set_options(traversal=mode)
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            spp_cpu = 4
            rays, dt, kind, tb = cpu_reference(cam, spp_cpu, threads)
            line["cpu_baseline"] = {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
                                    "sample": f"{W}x{H} x {spp_cpu} spp of the same frame ({rays} rays in {dt:.1f} s; reference BVH build {tb:.1f} s not included)"}
        if world == 1 and not args.no_cpu_baseline:
            line["legacy_gpu"] = legacy_gpu(s, cam)
        print(json.dumps(line), flush=True)
    if group is not None:
        group.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
