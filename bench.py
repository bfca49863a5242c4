#!/usr/bin/env python
"""bench.py -- headline benchmark of the render hot path (BASELINE.json: path samples/s and Mrays/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scene NAME] [--spp-step S]
                    [--shard samples|tiles] [--scaling weak|strong --spp-frame F]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one frame increment of the 1920x1080 teapot scene (BASELINE configs[1]): every GPU holds a replica of the scene and renders
its shard of the step's samples, then the partial estimators are combined on rank 0 INSIDE THE LIBRARY over NCCL
(paths_renderer_combine: reduce for sample shards, send/recv of owned tiles for tile shards).  Weak scaling (default): S
samples per pixel per GPU per step.  Strong scaling: F samples per pixel per step in total.  Rank 0 prints ONE JSON line.
`--impl reference` times the CPU restatement of the reference renderer (oracle/) on the host cores instead: the Rust reference itself
cannot be built here (no rustc/cargo).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
os.chdir(ROOT)

SCENES = {
    "teapot": ("scenes/teapot.yml", 1920, 1080),
    "teapot_light": ("scenes/teapot_light.yml", 1920, 1080),  # teapot.yml + the light of bunny.yml: NEE on the headline mesh
    "bunny": ("scenes/bunny.yml", 1920, 1080),
    "spheres_on_plane": ("scenes/spheres_on_plane.yml", 720, 480),
    "bokeh_demo": ("scenes/bokeh_demo.yml", 1920, 1080),
    "dragon": ("scenes/dragon.yml", 1920, 1080),        # synthetic 871k-triangle stand-in mesh (tools/make_assets.py)
    "environment": ("scenes/environment.yml", 1920, 1080),
    "doom": ("scenes/doom.yml", 1920, 1080),
}
ASSETS = {"dragon": ["dragon"], "environment": ["dragon", "hdri", "table"], "doom": ["doom"]}
SYNTHETIC = {"dragon": " (asset absent: synthetic 871k-triangle stand-in mesh, tools/make_assets.py)",
             "environment": " (assets absent: synthetic dragon, table and 4096x2048 HDRI stand-ins, tools/make_assets.py)",
             "doom": " (asset absent: synthetic vertex-coloured stand-in mesh, tools/make_assets.py)"}


def ensure_assets(scene):
    if scene in ASSETS:
        sys.path.insert(0, str(ROOT / "tools"))
        import make_assets

        make_assets.ensure(*ASSETS[scene])


def also_legs(args, world):
    """BASELINE.json quotes the metric on teapot AND dragon, and names environment + doom tile-sharded as config 5; teapot_light is the
    headline mesh with next-event estimation: the default (teapot) run of our arm adds short legs for them, outside the
    headline's timed region.  -> [(label, scene, shard)]"""
    if args.impl != "ours" or args.scene != "teapot" or args.also in ("", "none"):
        return []
    legs = []
    for x in args.also.split(","):
        name, _, shard = x.partition(":")
        if name in SCENES and name != args.scene:
            legs.append((x.replace(":", "_"), name, shard or args.shard))
    return legs


METRIC = "path_samples_per_sec"
UNIT = "samples/s"


def make_config(args, world, prims):
    """The `config` object: identical in both arms (the reference arm times a bounded sample of this workload and says which)."""
    yaml, W, H = SCENES[args.scene]
    if args.scaling == "strong":
        what = f"{args.spp_frame} spp per step in total (strong scaling)"
    else:
        what = f"{args.spp_step} spp per GPU per step"
    return {"workload": f"{yaml} at {W}x{H}, {what}" + SYNTHETIC.get(args.scene, ""), "scene": args.scene, "width": W, "height": H,
            "spp_per_step_per_gpu": args.spp_step if args.scaling == "weak" else args.spp_frame / max(world, 1), "prims": prims, "seed": args.seed,
            "shard": args.shard, "scaling": args.scaling,
            "l2": "per-pass path state (33M slots x 176 B = 5.8 GB per pass context) streams through HBM and exceeds the 126 MB L2; the scene tables "
                  "(BVH + triangles, ~1 MB for teapot) are L2-resident by design"}


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.device}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower() == "active":
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def reference_model_bytes(ctr):
    """bytes per BVH query Q and per sample, priced from the REFERENCE algorithm's own work counters (its AAC tree)."""
    q = max(ctr["extend_rays"] + ctr["shadow_rays"], 1)
    Q = 32 + 16 + (24 * ctr["box_tests"] + 40 * ctr["tri_tests"] + 16 * ctr["sphere_tests"]) / q
    per_sample = 64 + (208 * ctr["segments"] + Q * q) / max(ctr["samples"], 1)
    return Q, per_sample


def cpu_reference_leg(scene_yaml, W, H, budget_s, threads=0):
    """Times the oracle (CPU restatement of src/trace.rs + src/bvh.rs + src/worker.rs) on a bounded sample of the workload:
    whole frames of 1 sample per pixel until `budget_s` seconds of CPU work have been done."""
    import oracle

    osc, cam = oracle.load_scene(scene_yaml, width=W, height=H)
    ncores = threads or os.cpu_count() or 1
    total = oracle.Counters()
    done, t_total, s = 0, 0.0, 0
    while True:
        t0 = time.perf_counter()
        osc.render(cam, 1, s, s + 1, threads=ncores, counters=total)
        t_total += time.perf_counter() - t0
        s += 1
        done += W * H
        if t_total >= budget_s or s >= 64:
            break
    ctr = total.as_dict()
    return {"value": done / t_total, "unit": UNIT, "cores": ncores, "kind": "port",
            "sample": f"{s} spp of the full {W}x{H} frame ({done} camera paths, {t_total:.1f} s)",
            "mrays_per_sec": (ctr["extend_rays"] + ctr["shadow_rays"]) / t_total / 1e6}, ctr


def run_reference(args, rank):
    """--impl reference: the reference's own CPU path.  The Rust crate cannot be compiled in this image (no rustc/cargo, crates
    not vendored), so the timed program is the oracle port with all host threads; each step is one 1-spp frame of the workload."""
    if rank != 0:
        return
    import oracle

    yaml, W, H = SCENES[args.scene]
    osc, cam = oracle.load_scene(yaml, width=W, height=H)
    ncores = os.cpu_count() or 1
    for w in range(args.warmup):
        osc.render(cam, args.seed, w, w + 1, threads=ncores)
    ctr = oracle.Counters()
    t0 = time.perf_counter()
    for k in range(args.steps):
        osc.render(cam, args.seed, 100 + k, 101 + k, threads=ncores, counters=ctr)
    dt = time.perf_counter() - t0
    c = ctr.as_dict()
    value = c["samples"] / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": make_config(args, args.gpus, osc.num_prims),
        "mrays_per_sec": (c["extend_rays"] + c["shadow_rays"]) / dt / 1e6,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": ncores, "kind": "port",
                         "sample": f"each step = 1 spp of the full {W}x{H} frame of the config's workload ({args.steps} steps, {c['samples']} camera paths)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "Rust reference unbuildable here (no rustc/cargo); this is the C++ f64 restatement in oracle/, all host threads",
    }
    print(json.dumps(line))


def source_hash():
    """Hash of the kernel sources: the ncu-derived fields of `roofline` are printed only when profiles/*.json was captured on this code."""
    h = hashlib.sha256()
    for p in sorted((ROOT / "paths_b200" / "csrc").glob("*")):
        if p.suffix in (".cu", ".cuh", ".h", ".cpp"):
            h.update(p.name.encode())
            h.update(p.read_bytes())
    return h.hexdigest()[:16]


def capture_for(scene, fname):
    """profiles/traffic.json / issue.json entry of `scene` (tools/capture_roofline.sh), or (None, why) when it belongs to other code."""
    try:
        e = json.loads((ROOT / "profiles" / fname).read_text()).get(scene)
    except (OSError, ValueError):
        return None, f"profiles/{fname} unreadable"
    if not e or "classes" not in e:
        return None, f"no capture of {scene} in profiles/{fname}"
    if e.get("source_hash") != source_hash():
        return None, f"profiles/{fname} was captured on other kernel sources (head {e.get('git_head')}, hash {e.get('source_hash')}); re-run tools/capture_roofline.sh"
    return e, None


def rotation(yaw, pitch, roll):  # Matrix3::rotation (src/matrix.rs:30-35)
    import numpy as np

    cp, sp, cy, sy, cr, sr = math.cos(pitch), math.sin(pitch), math.cos(yaw), math.sin(yaw), math.cos(roll), math.sin(roll)
    rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rz = np.array([[cr, -sr, 0], [sr, cr, 0], [0, 0, 1]])
    return rx @ ry @ rz


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    yaml, W, H = SCENES[args.scene]
    legs = also_legs(args, world)

    # ---- CPU legs first, on rank 0 alone, BEFORE the process group exists: the other ranks wait in the rendezvous (CPU-side, no GPU
    # spin, no scene loading that would compete for the host cores).  The only use of oracle/ in this arm.
    cpu = cpu_t4 = ctr = None
    cpu_legs = {}
    if rank == 0 and args.cpu_budget > 0:
        cpu, ctr = cpu_reference_leg(yaml, W, H, args.cpu_budget)
        cpu_t4, _ = cpu_reference_leg(yaml, W, H, min(args.cpu_budget, 6.0), threads=4)  # the reference's worker count (src/main.rs:87)
        cpu_t4["note"] = "Renderer::new(camera, scene, 4): the reference's compile-time worker count (src/main.rs:87)"
        done = set()
        for label, name, _ in legs:
            if name not in done:
                y2, W2, H2 = SCENES[name]
                cpu_legs[name], _ = cpu_reference_leg(y2, W2, H2, min(args.cpu_budget, 6.0))
                done.add(name)

    import paths_b200 as pb

    if not torch.cuda.is_available() or pb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; paths_b200 has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    S = args.spp_step
    spp_call = (S * world) if args.scaling == "weak" else args.spp_frame  # samples per pixel one step adds to the frame, all ranks together

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def make_renderer(scene, timed, shard):
        """A renderer of this rank's shard, joined to a fresh NCCL communicator of the library (unique id broadcast through torch)."""
        r = pb.Renderer(scene.camera, scene, device=local_rank, precision=args.precision, seed=args.seed, rank=rank, world=world,
                        shard_mode=pb.SHARD_TILES if shard == "tiles" else pb.SHARD_SAMPLES, time_kernels=timed)
        if world > 1:
            box = [pb.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0)
            r.attach_comm(box[0])
        return r

    def step(r, spp):
        # sample shards: this rank renders its share of the `spp` samples of every pixel; tile shards: all `spp` samples of its own
        # 1/world of the pixels; then the combine on rank 0, enqueued on the renderer's stream inside the library
        r.render_samples(spp, wait=False)
        if world > 1:
            r.combine(0, wait=False)

    def timed_region(r, steps, warmup, spp, sample_clocks):
        """W warm-up steps, then exactly `steps` steps between two CUDA events recorded on the renderer's own stream."""
        ext = torch.cuda.ExternalStream(r.cuda_stream(), device=local_rank)
        with torch.cuda.stream(ext):
            for _ in range(warmup):
                step(r, spp)
            r.drain_result_queue()
            barrier()
            st0 = r.stats()
            clocks = ClockSampler(local_rank) if sample_clocks else None
            if clocks:
                clocks.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            for _ in range(steps):
                step(r, spp)
            e1.record(ext)
            r.drain_result_queue()
            barrier()
            ms = e0.elapsed_time(e1)
            clk = clocks.stop() if clocks else None
        st1 = r.stats()
        ranks_ms = [ms]
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            allt = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allt, t)
            ranks_ms = [float(x.item()) for x in allt]
        delta = {k: st1[k] - st0[k] for k in st1 if k not in ("extend_kernel", "camera_kernel", "shade_kernel", "pass_contexts")}
        kernels = {k: st1[k] for k in ("extend_kernel", "camera_kernel", "shade_kernel", "pass_contexts")}
        return max(ranks_ms), ranks_ms, delta, clk, kernels

    def totals(d, keys):
        t = torch.tensor([d[k] for k in keys], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(x) for x in t.tolist()]

    scene = pb.Scene.load(yaml, width=W, height=H)
    info = scene.info

    # ---- headline: inputs resident in HBM, device-timed, max over ranks
    r = make_renderer(scene, False, args.shard)
    ms, ranks_ms, d, clocks, kernels = timed_region(r, args.steps, args.warmup, spp_call, True)
    tot_samples, tot_ext, tot_shadow, tot_launch = totals(d, ("samples", "extend_rays", "shadow_rays", "kernel_launches"))
    value = tot_samples / (ms * 1e-3)
    overflows = d["stack_overflows"]
    r.shutdown()

    # ---- per-kernel split: the same steps again with CUDA events around every kernel (one pass context, no overlap)
    rt = make_renderer(scene, True, args.shard)
    ms_t, _, dt_, _, _ = timed_region(rt, args.steps, args.warmup, spp_call, False)
    rt.shutdown()

    # ---- end to end through the C ABI with host buffers, on all N GPUs: camera in (new epoch) on every rank, each rank renders
    # its shard, combine on rank 0 inside the library, RGB24 frame out to pinned host memory on rank 0, every step
    re_ = make_renderer(scene, False, args.shard)
    cam = scene.camera
    rot = rotation(cam.yaw, cam.pitch, cam.roll)
    loc = np.array(list(cam.location))
    out = torch.empty((H, W, 3), dtype=torch.uint8, pin_memory=True).numpy() if rank == 0 else None

    def e2e_step():
        re_.set_camera(loc, rot)                      # host -> device: camera (new epoch, estimator cleared)
        step(re_, spp_call)
        re_.drain_result_queue()
        if rank == 0:
            pb.check(pb.lib().paths_renderer_image_rgb24(re_._h, out.ctypes.data))  # device -> host: tonemapped combined frame

    e2e_steps = max(3, min(args.steps, 20))
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_dt = float(t.item())
    e2e_value = e2e_steps * spp_call * W * H / e2e_dt
    re_.shutdown()

    # ---- N GPUs = 1 GPU: the frame the ranks rendered and combined over NCCL against the same frame rendered unsharded on rank 0
    # (tile shards: bit-identical; sample shards: equal up to the order of the f32 additions)
    multi_gpu_check = None
    if world > 1 and not args.no_check:
        multi_gpu_check = {}
        spp_chk = min(spp_call, 8 * world)
        for shard in ("tiles", "samples"):
            rc = make_renderer(scene, False, shard)
            rc.render_samples(spp_chk // 2, wait=False)
            rc.combine(0, wait=False)
            rc.render_samples(spp_chk - spp_chk // 2, wait=False)  # a second increment: the combine must track the cumulative estimator
            rc.combine(0, wait=True)
            got = rc.sums() if rank == 0 else None
            rc.shutdown()
            if rank == 0:
                one = pb.Renderer(scene.camera, scene, device=local_rank, precision=args.precision, seed=args.seed)
                one.render_samples(spp_chk // 2)
                one.render_samples(spp_chk - spp_chk // 2)
                ref = one.sums()
                one.shutdown()
                counts_equal = bool(np.array_equal(got[..., 3], ref[..., 3]))
                diff = np.abs(got[..., :3] - ref[..., :3])
                rel = float((diff / (np.abs(ref[..., :3]) + 1e-6)).max())
                res = {"spp": spp_chk, "counts_equal": counts_equal, "bit_identical": bool(np.array_equal(got, ref)), "max_rel_diff": rel}
                res["pass"] = counts_equal and (res["bit_identical"] if shard == "tiles" else rel <= 2e-6 + 1e-6)
                multi_gpu_check[shard] = res
            barrier()
        if rank == 0:
            multi_gpu_check["pass"] = all(v["pass"] for v in multi_gpu_check.values() if isinstance(v, dict))
            multi_gpu_check["how"] = ("N-rank frame combined by paths_renderer_combine (NCCL inside libpaths_b200.so) vs the same frame rendered "
                                      "by one renderer on rank 0; tiles must be bit-identical, sample shards within rtol 2e-6")

    # ---- strong scaling of the headline scene: a fixed frame of `spp_frame` samples per pixel split over the N ranks (256 spp),
    # combine inside the timed region; at N = 1 the same frame on one GPU, so that the driver's 1 / 2 / 4 / 8 runs give the curve
    strong = None
    if args.scaling == "weak" and not args.no_strong:
        rs = make_renderer(scene, False, args.shard)
        ks = 3
        ms_s, ranks_s, d_s, _, _ = timed_region(rs, ks, 3, args.spp_frame, False)
        rs.shutdown()
        smp_s, = totals(d_s, ("samples",))
        strong = {"scaling": "strong", "spp_frame": args.spp_frame, "value": smp_s / (ms_s * 1e-3), "unit": UNIT, "ms_per_frame": ms_s / ks, "steps": ks,
                  "warmup": 3, "n_gpus": world, "rank_ms_max": max(ranks_s) / ks, "rank_ms_mean": sum(ranks_s) / len(ranks_s) / ks,
                  "what": f"{yaml} at {W}x{H}: one {args.spp_frame}-spp frame per step in total, {args.shard} shards, combine on rank 0 inside the timed region"}

    # ---- BASELINE's other configs (dragon; environment and doom tile-sharded): short device-timed legs of the same step
    others = {}
    for label, name, shard in legs:
        ensure_assets(name)
        yaml2, W2, H2 = SCENES[name]
        scene2 = pb.Scene.load(yaml2, width=W2, height=H2)
        r2 = make_renderer(scene2, False, shard)
        k2 = max(3, min(args.steps, 5))
        ms2, ranks2, d2, _, kern2 = timed_region(r2, k2, 3, S * world, False)
        r2.shutdown()
        smp2, ext2, sh2 = totals(d2, ("samples", "extend_rays", "shadow_rays"))
        others[label] = {"value": smp2 / (ms2 * 1e-3), "unit": UNIT, "mrays_per_sec": (ext2 + sh2) / (ms2 * 1e-3) / 1e6, "ms_per_step": ms2 / k2,
                         "steps": k2, "warmup": 3, "n_gpus": world, "spp_per_step_per_gpu": S, "prims": scene2.info["n_prims"], "shard": shard,
                         "rank_ms_max": max(ranks2) / k2, "rank_ms_mean": sum(ranks2) / len(ranks2) / k2,
                         "precision": scene2.info["auto_precision"], "extend_kernel": kern2["extend_kernel"],
                         "workload": f"{yaml2} at {W2}x{H2}, {S} spp per GPU per step" + SYNTHETIC.get(name, "")}
        if name in cpu_legs:
            others[label]["cpu_baseline"] = cpu_legs[name]
        del scene2

    line = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except (OSError, ValueError):
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        # ---- roofline.  Whole scenes sit in L1/L2 (teapot tree < 1 MB), so what HBM must carry is the wavefront's own stream traffic:
        # per sample A = 64 B (new path record + estimator RMW), per path segment S = 208 B (shade: ray + hit + material in, next ray +
        # shadow record + path state out), per BVH query 48 B (ray in, hit / visibility out) -- SURVEY 8(d)'s formula without its tree terms.
        n_samples, n_ext, n_sh = max(d["samples"], 1), d["extend_rays"], d["shadow_rays"]
        stream_bytes_per_sample = 64 + (208 * n_ext + 48 * (n_ext + n_sh)) / n_samples
        rank_ms_step = ranks_ms[0] / args.steps
        step_gbs = stream_bytes_per_sample * n_samples / args.steps / (rank_ms_step * 1e-3) / 1e9
        split = {k: dt_[k] for k in ("ms_raygen", "ms_extend", "ms_shade", "ms_shadow", "ms_accumulate")}
        dom = max(("ms_extend", "ms_shade", "ms_shadow"), key=lambda k: split[k])
        dom_ms = split[dom]
        n_dom_launch = max(dt_["n_extend_launches"] if dom == "ms_extend" else dt_["n_shadow_launches"] if dom == "ms_shadow" else dt_["n_extend_launches"], 1)
        dom_bytes = {"ms_extend": 48 * dt_["extend_rays"], "ms_shadow": 48 * dt_["shadow_rays"], "ms_shade": 208 * dt_["extend_rays"]}[dom]
        dom_gbs = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
        roof = {"bound": "hbm", "achieved": step_gbs, "peak": peak, "unit": "GB/s", "frac": step_gbs / peak,
                "what": "whole step on rank 0: compulsory stream bytes of the wavefront layout (64 B per sample + 208 B per path segment + 48 B per BVH "
                        "query, from this run's own counters) over the device-timed step; the scene tables are cache-resident and priced at zero",
                "algorithmic_bytes_per_sample": stream_bytes_per_sample, "samples_per_step": n_samples / args.steps,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst)" if peaks else "fallback 6650 (B200_PROFILING.md)",
                "split_ms": split, "kernel_share_of_step": dom_ms / max(dt_["ms_total"], 1e-9),
                "kernel": {"ms_extend": "closest-hit kernels: k_camera_extend (raygen + camera segment) + " + {3: "k_extend_refill<1,16,8>", 16: "k_extend4s<1,16,16,12>", 0: "k_extend"}.get(kernels["extend_kernel"], "?") + " (bounces)",
                           "ms_shade": "k_shade", "ms_shadow": "visibility kernels"}[dom],
                "kernel_launches": int(n_dom_launch), "kernel_avg_launch_ms": dom_ms / n_dom_launch,
                "kernel_achieved": dom_gbs, "kernel_frac": dom_gbs / peak,
                "kernel_algorithmic_bytes_per_launch": dom_bytes / n_dom_launch,
                "kernels_in_use": kernels, "stack_overflows": int(overflows)}
        # the figure round 1 printed as `frac`: the REFERENCE algorithm's traversal work (its AAC tree, 36 box tests per ray on teapot)
        # priced as HBM traffic -- kept for the record under its own name, not a physical fraction
        if ctr is not None:
            Q, per_sample_ref = reference_model_bytes(ctr)
            ext_ms = split["ms_extend"]
            roof["frac_reference_model"] = (Q * dt_["extend_rays"]) / (ext_ms * 1e-3) / 1e9 / peak if ext_ms > 0 else None
            roof["reference_model"] = {"bytes_per_ray": Q, "bytes_per_sample": per_sample_ref,
                                       "note": "SURVEY 8(d) prices the reference tree's box / triangle tests as HBM bytes; the production tree does "
                                               "far less work and stays in cache, so this exceeds 1 and binds nothing"}
        # ncu-derived fields, only from a capture of THIS code (tools/capture_roofline.sh stamps git head + source hash)
        cls = {"ms_extend": "extend", "ms_shade": "shade", "ms_shadow": "shadow"}[dom]
        tj, why = capture_for(args.scene, "traffic.json")
        if tj and world == 1:
            c = tj["classes"]
            pool = [c[k] for k in (("camera", "extend") if cls == "extend" else (cls,)) if k in c]
            n_l = sum(x["launches"] for x in pool)
            byt = sum(x["dram_bytes"] for x in pool)
            roof["traffic"] = byt / max(n_l, 1)
            scale = (n_samples / args.steps) / (1920 * 1080 * 16)  # the capture is one 16-spp pass; a step is this many of them
            kernel_gbs = byt * scale / (dom_ms / args.steps * 1e-3) / 1e9 if dom_ms > 0 else None  # that class's DRAM bytes per step over its time per step
            roof["dram_gbs_measured"] = kernel_gbs
            roof["dram_frac_measured"] = kernel_gbs / peak if kernel_gbs is not None else None
            roof["dram_gbs_measured_step"] = sum(x["dram_bytes"] for x in c.values()) * scale / (rank_ms_step * 1e-3) / 1e9
            roof["dram_frac_measured_step"] = roof["dram_gbs_measured_step"] / peak
            roof["traffic_source"] = f"profiles/traffic.json[{args.scene}]: ncu dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel class, git head {tj['git_head']}, source hash {tj['source_hash']}"
        else:
            roof["traffic"] = None
            roof["traffic_source"] = why if world == 1 else "single-GPU capture; not applied to a multi-GPU line"
        ij, why_i = capture_for(args.scene, "issue.json")
        if ij:
            c = ij["classes"]
            roof["issue"] = {k: {"issue_active_pct": v["issue_active_pct"], "lanes_per_instruction": v["lanes_per_instruction"],
                                 "frac_of_issue_roofline": v["issue_x_lanes"], "share_of_pass": v["share_of_pass"]} for k, v in c.items() if k != "other"}
            roof["issue"]["source"] = f"profiles/issue.json[{args.scene}]: ncu smsp__issue_active x smsp__thread_inst_executed_per_inst_executed / 32, time-weighted per kernel class, git head {ij['git_head']}, source hash {ij['source_hash']}"
            top = max((k for k in c if k != "other"), key=lambda k: c[k]["ms"])
            roof["limiter"] = (f"{top}: {c[top]['issue_active_pct']:.0f} % of issue slots busy with {c[top]['lanes_per_instruction']:.1f} of 32 lanes active "
                               f"= {c[top]['issue_x_lanes']:.2f} of the SIMT issue roofline; DRAM {c[top]['dram_gbs'] / peak:.2f} of measured HBM peak during that kernel")
        else:
            roof["issue"] = None
            roof["limiter"] = why_i
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f32" if (args.precision or info["auto_precision"]) == 32 else "f64", "data": "synthetic",
            "config": make_config(args, world, info["n_prims"]),
            "mrays_per_sec": (tot_ext + tot_shadow) / (ms * 1e-3) / 1e6,
            "rays_per_sample": (tot_ext + tot_shadow) / max(tot_samples, 1),
            "rank_ms_per_step": {"max": max(ranks_ms) / args.steps, "mean": sum(ranks_ms) / len(ranks_ms) / args.steps, "min": min(ranks_ms) / args.steps},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": (12 * 8 + 208) * world, "d2h_bytes_per_step": W * H * 3,
                    "what": f"set_camera (new epoch) + render + RGB24 frame to pinned host memory through the C ABI on {world} GPU(s)"
                            + (", combine on rank 0 inside the library (NCCL) inside the timed region" if world > 1 else "")},
            "gpu_launches": int(tot_launch),
            "combine": None if world == 1 else ("paths_renderer_combine inside libpaths_b200.so: " + ("ncclSend/ncclRecv of owned 32x32 tiles" if args.shard == "tiles" else "ncclReduce(sum) of the float4 estimator") + f", {pb.nccl_info()}"),
            "roofline": roof,
            "cpu_baseline": cpu,
            "cpu_baseline_t4": cpu_t4,
        }
        if multi_gpu_check is not None:
            line["multi_gpu_check"] = multi_gpu_check
        if strong is not None:
            line["strong_scaling"] = strong
        if others:
            line["other_scenes"] = others
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="teapot", choices=sorted(SCENES))
    ap.add_argument("--spp-step", type=int, default=64, help="weak scaling: samples per pixel per GPU per step")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--spp-frame", type=int, default=256, help="strong scaling: samples per pixel per step in total (SURVEY C2/C4: 256)")
    ap.add_argument("--shard", default="samples", choices=["samples", "tiles"],
                    help="N > 1: each GPU renders a sample range of every pixel (default) or all samples of its 32x32 tiles")
    ap.add_argument("--precision", type=int, default=0, choices=[0, 32, 48, 64])
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--no-strong", action="store_true", help="skip the short strong-scaling leg of the default run")
    ap.add_argument("--also", default="dragon,teapot_light,environment:tiles,doom:tiles",
                    help="comma-separated extra legs `scene[:shard]` timed briefly after the teapot headline (none = skip)")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--no-check", action="store_true", help="skip the N-GPU = 1-GPU frame check (N > 1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if rank == 0:
        ensure_assets(args.scene)
        for _, extra, _ in also_legs(args, world):
            ensure_assets(extra)
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit(f"bench.py: --gpus {args.gpus} needs `python -m torch.distributed.run --nproc-per-node {args.gpus} bench.py ...`")
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
