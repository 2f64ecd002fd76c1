#!/usr/bin/env python
"""bench.py — one JSON line per run (contract in the task statement, tier reading).

A step = one pass of the hot path (raygen -> [extend -> shade] x depth -> accumulate) over one frame of the
headline workload, BASELINE.json configs[1]: procedural 262 144-triangle atrium, 1024x1024, 64 spp, max depth 5,
one B200.  With N GPUs every rank renders its own 64-sample range of the same frame (disjoint Philox sample
indices, replicated BVH) and the f32 accumulators are summed on rank 0 with one NCCL reduce: weak scaling.

  value      counted Mrays/s (one ray = one extend invocation) of the whole job, scene + BVH resident in HBM
  e2e        the same through the C-ABI with HOST buffers: cupt_set_scene + cupt_set_bvh (H2D of the scene arrays
             and the BVH image), cupt_render, cupt_read_accum (D2H of the image) inside the timed region
  roofline   extend kernel: algorithmic bytes (node visits x 80 B + triangle tests x 48 B + 44 B of ray/hit/queue
             records per ray, counted by a stats pass of the same step) / its CUDA-event time, against the measured
             HBM copy bandwidth
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


def ncu_traffic():
    """per-launch DRAM bytes of the extend kernel from the committed ncu --set full summary, if there is one"""
    p = os.path.join(ROOT, "profiles", "extend_ncu_summary.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def cpu_reference(scene_handle, cam, spp, threads):
    """the reference's integrator on host cores: (rays, seconds, kind).  oracle/_ref when it was built, else the port."""
    import oraclebind
    import refbind
    v, t, m, l = scene_handle.arrays()
    area = scene_handle.total_light_area
    if refbind.available():
        refbind.reset()
        refbind.set_arrays(v, t, m, l, area)
        devnull = os.open(os.devnull, os.O_WRONLY)
        saved = os.dup(1)
        os.dup2(devnull, 1)   # buildBVH() printf()s (BVH.h:379-380,445,464)
        try:
            t0 = time.perf_counter(); refbind.build_bvh(); tb = time.perf_counter() - t0
        finally:
            os.dup2(saved, 1); os.close(devnull); os.close(saved)
        t0 = time.perf_counter()
        _, _, rays = refbind.render(cam, integrator=2, num_bounces=DEPTH, rng=1, seed=SEED, first_sample=0, num_samples=spp, nthreads=threads)
        return rays, time.perf_counter() - t0, "reference", tb
    n2, d2 = scene_handle.build_bvh2()
    osc = oraclebind.OracleScene(v, t, m, l, area, n2)
    t0 = time.perf_counter()
    _, _, rays = osc.render_pt(cam, DEPTH, SEED, 0, spp)
    return rays, time.perf_counter() - t0, "port", 0.0


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
    sc, s, cam = build_scene()
    threads = os.cpu_count() or 1
    spp = 1   # bounded sample per step: 1024x1024 x 1 spp of the same frame
    times, rays_total = [], 0
    kind = "port"
    for i in range(args.warmup + args.steps):
        rays, dt, kind, _ = cpu_reference(s, cam, spp, threads)
        if i >= args.warmup:
            times.append(dt); rays_total += rays
    total = sum(times)
    val = rays_total / total / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64/f32 (reference: f32 geometry, f64 colour)", "data": "synthetic",
            "config": {"workload": WORKLOAD, "step": f"{W}x{H} x {spp} spp of the frame on host cores"},
            "msamples_per_s": W * H * spp * args.steps / total / 1e6,
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": kind,
                             "sample": f"{W}x{H} x {spp} spp per step, radianceAlongSingleStep2 depth {DEPTH}, keyed Philox"},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    h2d = nv * 12 + nt * 28 + nt * 16 + nm * 32 + (nl + 1) * 48 + 32
    if args.bvh == "host":
        h2d += info["num_nodes"] * 80 + info["num_tri_records"] * 48      # the device builder uploads no tree
    if mode != api.TRAVERSAL_WIDE:
        h2d += len(n2) * 32 * 2 + nt * 4                                   # reference-layout tree, its parent-linked copy, leaf parents
    d2h = W * H * 12

    def e2e_step():
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

    if rank == 0:
        peak, peak_src = measured_peak()
        n_ext = max(1, st["extend_launches"])
        ext_ms = st["ms_extend"]
        bytes_step = nodes_step * 80 + tris_step * 48 + rays_step * 44
        achieved = bytes_step * args.steps / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None
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
            "e2e": {"value": e2e_val, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n_e2e},
            "gpu_launches": int(st["kernel_launches"]),
            "stage_ms_per_step": {k: st["ms_" + k] / args.steps for k in ("raygen", "extend", "shade", "accum")},
            "roofline": {"bound": "hbm", "kernel": "k_extend<0,false,*>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": (achieved / peak) if achieved else None, "peak_source": peak_src, "traffic": ncu_traffic(),
                         "algorithmic_bytes_per_launch": bytes_step / (n_ext / args.steps) if n_ext else None,
                         "launch_ms": ext_ms / n_ext, "launches": int(n_ext),
                         "bytes_per_ray": bytes_step / rays_step, "nodes_per_ray": nodes_step / rays_step, "tris_per_ray": tris_step / rays_step},
        }
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
            rays, dt, kind, tb = cpu_reference(s, cam, spp_cpu, threads)
            line["cpu_baseline"] = {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
                                    "sample": f"{W}x{H} x {spp_cpu} spp of the same frame ({rays} rays in {dt:.1f} s; reference BVH build {tb:.1f} s not included)"}
        if world == 1 and not args.no_cpu_baseline:
            line["legacy_gpu"] = legacy_gpu(s, cam)
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
