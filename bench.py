#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json metric: Mrays/s + samples/s, 1080p Sponza-like
262 144-triangle atrium, depth 5, NEE + Russian roulette).

  python bench.py [--gpus N] [--steps K] [--warmup W]          our CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference ...                         the reference's algorithm on the host cores (CPU oracle)

A "step" is one frame = one path (sample) per pixel of the 1920x1080 image, i.e. one pass of
raygen -> 5 x (extend, shade, shadow) over 2 073 600 paths. One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WIDTH, HEIGHT, MAX_DEPTH, TRIS = 1920, 1080, 5, 262144
METRIC = "Mrays/s (closest-hit + shadow rays), 1080p atrium-262k path tracing, depth 5, NEE + RR"

# BASELINE.json configs -> workloads. The default (and the driver's) run is configs[1] = "atrium"; the others are
# selected with --workload and recorded under profiles/ by tools/run_configs.py.
WORKLOADS = {
    "atrium": dict(kind="pt", width=1920, height=1080, depth=5, model=0, tris=262144, seed=2, aperture=0.0, glossy=False,
                   desc="atrium-262144 1080p depth-5 NEE+RR (BASELINE.json configs[1], procedural stand-in for Sponza)"),
    "bathroom": dict(kind="pt", width=1920, height=1080, depth=5, model=1, tris=1000000, seed=3, aperture=0.5, glossy=True,
                     desc="glossy interior ~1.0 M triangles, Beckmann microfacet + thin-lens DOF (lens radius 0.25), 1080p depth-5 "
                          "(BASELINE.json configs[2], procedural stand-in for Salle de Bain)"),
    "4k": dict(kind="pt", width=4096, height=2160, depth=5, model=0, tris=262144, seed=2, aperture=0.0, glossy=False,
               desc="atrium-262144 4096x2160 depth-5 NEE+RR, samples split over the GPUs + NCCL reduce (BASELINE.json configs[4])"),
    "micro": dict(kind="rays", n_rays=1 << 24, desc="16 777 216 random rays (origins uniform in the AABB, directions uniform on the "
                  "sphere, unsorted) vs 8 nested displaced shells = 10 010 896 triangles: closest-hit + shadow any-hit "
                  "(BASELINE.json configs[3])"),
    "bunny": dict(kind="primary", width=1280, height=720,
                  desc="bunny-like 69 196-triangle mesh: SAH/SBVH build + 1 spp primary closest-hit at 1280x720 (BASELINE.json configs[0])"),
}
# algorithmic bytes of the HBM data layout: ray in 32 B, hit out 16 B, node pair 64 B, leaf triangle 48 B
B_RAY_IN, B_HIT_OUT, B_PAIR, B_TRI, B_FLAG_OUT = 32, 16, 64, 40, 4


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.02)
        except Exception as e:      # NVML missing: report it rather than hide it
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def build_scene(wl=None):
    from gpupathtracer_b200 import Scene, scenes
    wl = wl or WORKLOADS["atrium"]
    data = scenes.atrium(target_tris=wl["tris"], seed=wl["seed"], aperture=wl["aperture"], glossy=wl["glossy"])
    return Scene(data)


def make_camera(scene, w, h, aspect=None):
    from gpupathtracer_b200 import Camera
    c = scene.data.camera
    cam = Camera(aspect or (w / h), c["fov"], c["focus"], c["aperture"])
    cam.SetPosition(c["pos"]); cam.SetRotation(c["rot"]); cam.GenerateImagePlane()
    return cam


def traffic_for(workload, kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        return json.load(f).get(workload, {}).get(kernel)


def cpu_reference_run(scene, steps, warmup, w, h, wl, threads=0):
    """The reference's algorithm (CPU restatement, oracle/) on a bounded sample of the workload: the same scene,
    camera, depth and estimator at w x h pixels, 1 frame per step. Returns (Mrays/s, samples/s, ms/step, counters)."""
    import oracle
    from gpupathtracer_b200 import make_rng_table
    o = oracle.Oracle(scene)
    cam = make_camera(scene, w, h, aspect=wl["width"] / wl["height"]).as_array()
    rng = make_rng_table(w * h, 1234)
    acc = np.zeros((h, w, 4), np.float32)
    kw = dict(n_frames=1, max_depth=wl["depth"], model=wl["model"], n_threads=threads)
    for _ in range(warmup):
        o.render(cam, w, h, rng, acc, **kw)
    tot = {"closest_rays": 0, "shadow_rays": 0, "node_pairs": 0, "tri_tests": 0, "shadow_node_pairs": 0, "shadow_tri_tests": 0}
    t0 = time.perf_counter()
    for _ in range(steps):
        c = o.render(cam, w, h, rng, acc, **kw)
        for k in tot:
            tot[k] += c[k]
    dt = time.perf_counter() - t0
    rays = tot["closest_rays"] + tot["shadow_rays"]
    info = {"kind": "port", "what": "the GLSL megakernel restated in C++ (oracle/pt_oracle.cpp), all host threads"}
    if oracle.have_glsl() and wl["model"] in (0, 1):
        # The reference's OWN main() (Iterative.comp compiled as C++, oracle/_ref/libref_glsl.so) on the same frames: one
        # process per core sharing the ray counter. Same pixels, same RNG table, hence the same rays as counted above.
        try:
            procs = threads or (os.cpu_count() or 1)
            rng2 = make_rng_table(w * h, 1234)
            acc2 = np.zeros((h, w, 4), np.float32)
            if warmup:
                o.ref_glsl_render_mp(cam, w, h, rng2, acc2, n_frames=warmup, max_depth=wl["depth"], model=wl["model"], n_procs=procs)
            t0 = time.perf_counter()
            o.ref_glsl_render_mp(cam, w, h, rng2, acc2, n_frames=steps, max_depth=wl["depth"], model=wl["model"], n_procs=procs)
            dt2 = time.perf_counter() - t0
            eq = (acc2.view(np.uint32) == acc.view(np.uint32)).all(axis=2)
            nan_px = np.isnan(acc2).any(axis=2)                       # Beckmann only: NaN through GLSL max(), DESIGN.md section 2
            same = bool(np.array_equal(rng2, rng)) and bool((eq | nan_px).all())
            info = {"kind": "reference", "what": "the reference's own main() (Iterative.comp:214-307 and callees compiled as C++, "
                    "oracle/_ref/libref_glsl.so; reference GLSL via a GL stack: unavailable on this box), one process per host "
                    "core sharing the ray counter", "port_value": rays / dt / 1e6, "same_bits_as_port": same,
                    "nan_pixels_in_glm_compile": int(nan_px.sum())}
            dt = dt2
        except Exception as e:                                       # never lose the bench line to the baseline leg
            info["reference_leg_error"] = repr(e)
    return rays / dt / 1e6, steps * w * h / dt, dt / steps * 1e3, tot, info


def metric_name(name, wl):
    if name == "atrium":
        return METRIC
    if wl["kind"] == "pt":
        return f"Mrays/s (closest-hit + shadow rays), {name} path tracing {wl['width']}x{wl['height']} depth {wl['depth']}, NEE + RR"
    if wl["kind"] == "rays":
        return "Mrays/s (closest-hit + shadow any-hit), 16M random rays vs 10M-triangle mesh"
    return "Mrays/s (primary closest-hit), bunny-69k 1280x720 1 spp (+ BVH build ms)"


CPU_SAMPLE = (480, 272)


def micro_scene_and_rays(n_rays, seed):
    from gpupathtracer_b200 import Scene, scenes
    scene = Scene(scenes.shells())
    lo, hi = scene.bounds()
    closest = scenes.random_rays(n_rays, lo, hi, seed=seed)
    anyr = closest.copy()
    diag = float(np.linalg.norm(hi - lo))
    anyr[:, 7] = np.random.default_rng(seed + 1).uniform(0, diag, size=n_rays).astype(np.float32)
    return scene, closest, anyr


def bytes_per_ray(c, n, out_bytes):
    """B_ray = 32 (ray in) + out + pairs/ray * 64 + tris/ray * 40 on the tree the kernel traverses."""
    pairs, tris = c["node_pairs"] / n, c["tri_tests"] / n
    return B_RAY_IN + out_bytes + pairs * B_PAIR + tris * B_TRI, pairs, tris


def _cpu_tracer(scene):
    """The CPU traversal the reference arm times: the reference's own IfIfClosestHit / IfIfAnyHit shader source compiled as C++
    (oracle/_ref/libref_glsl.so, a thread per core) when it was built, else the oracle's restatement."""
    import oracle
    if oracle.have_glsl():
        return (oracle.RefGlsl(scene.nodes, scene.refs, scene.tris),
                "the reference's own traversal shader source (common/BVH.glsl + Geometry.glsl compiled as C++), all host threads")
    return oracle.Oracle(scene), "the traversal restated in C++ (oracle/pt_oracle.cpp), all host threads"


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores (oracle/, all threads), bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    from gpupathtracer_b200 import build_bvh_host
    wl = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    kind = "port"
    base = {"impl": "reference", "metric": metric_name(args.workload, wl), "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic"}

    def ref_tree(scene):
        if oracle.have_ref():
            t0 = time.perf_counter()
            nodes, refs = oracle.ref_build_sbvh(scene.tris_abs)      # the reference's own builder, single-threaded as shipped
            return nodes, refs, "reference BuildBinnedSpatial (oracle/_ref)", (time.perf_counter() - t0) * 1e3
        t0 = time.perf_counter()
        nodes, refs, _ = build_bvh_host(scene.tris)
        return nodes, refs, "libb200pt host builder (oracle/_ref absent)", (time.perf_counter() - t0) * 1e3

    if wl["kind"] == "pt":
        scene = build_scene(wl)
        nodes, refs, tree, _ = ref_tree(scene)
        scene.set_bvh(nodes, refs)
        w, h = CPU_SAMPLE
        mrays, sps, ms, tot, info = cpu_reference_run(scene, args.steps, args.warmup, w, h, wl)
        sample = f"{w}x{h} px of the same camera/scene, 1 frame per step; {info['what']}"
        kind = info["kind"]
        extra = {"samples_per_s": sps, "reference_leg": {k: v for k, v in info.items() if k not in ("kind", "what")}}
    elif wl["kind"] == "rays":
        n = 1 << 20
        scene, closest, anyr = micro_scene_and_rays(n, 0xB200)
        nodes, refs, _ = build_bvh_host(scene.tris)                  # the reference builder needs minutes at 10 M triangles
        tree = "libb200pt host builder"
        scene.set_bvh(nodes, refs)
        o, what = _cpu_tracer(scene)
        kind = "reference" if oracle.have_glsl() else "port"
        for _ in range(args.warmup):
            o.trace_closest(closest[:1 << 16])
        t0 = time.perf_counter()
        for _ in range(args.steps):
            o.trace_closest(closest); o.trace_any(anyr)
        dt = time.perf_counter() - t0
        mrays, ms = 2 * n * args.steps / dt / 1e6, dt / args.steps * 1e3
        sample = f"first {n} closest-hit + {n} any-hit rays of the same batch per step; {what}"
        extra = {}
    else:
        from gpupathtracer_b200 import Scene, scenes
        scene = Scene(scenes.bunny_like(seed=1))
        nodes, refs, tree, build_ms = ref_tree(scene)
        scene.set_bvh(nodes, refs)
        o = oracle.Oracle(scene)
        w, h = wl["width"], wl["height"]
        from gpupathtracer_b200 import make_rng_table
        rng = make_rng_table(w * h, 1)
        cam = make_camera(scene, w, h).as_array()
        rays = np.zeros((w * h, 8), np.float32)
        pr = o.raygen(cam, w, h, rng, 0, w * h)
        rays[:, 0:3] = pr[:, 2:5]; rays[:, 4:7] = pr[:, 5:8]; rays[:, 7] = 1e20
        o, what = _cpu_tracer(scene)
        kind = "reference" if oracle.have_glsl() else "port"
        for _ in range(args.warmup):
            o.trace_closest(rays[:1 << 16])
        t0 = time.perf_counter()
        for _ in range(args.steps):
            o.trace_closest(rays)
        dt = time.perf_counter() - t0
        mrays, ms = w * h * args.steps / dt / 1e6, dt / args.steps * 1e3
        sample = f"all 921 600 primary rays per step; {what}"
        extra = {"bvh_build_ms": build_ms, "bvh_builder": tree}
    line = dict(base, value=mrays, ms_per_step=ms, config={"workload": wl["desc"], "tree": tree},
                cpu_baseline={"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
                e2e={"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, **extra)
    print(json.dumps(line))


class Dist:
    """torch.distributed plumbing shared by the workloads (one process per GPU, NCCL): rendezvous, barriers, max / sum of
    the per-rank timings. The data path's own exchange -- the reduce of the accumulation buffers -- runs inside the library
    (csrc/pt_comm.cu) over its own NCCL communicator, whose id travels through this process group."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self, r=None):
        if r is not None:
            r.CommSynchronize() if self.world > 1 else r.Synchronize()
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def max_sum(self, values):
        """values: list of floats; element 0 is reduced with MAX (time), the rest with SUM (work)."""
        t = self.torch.tensor(values, dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            tmax = t.clone(); self.dist.all_reduce(tmax, op=self.dist.ReduceOp.MAX)
            tsum = t.clone(); self.dist.all_reduce(tsum, op=self.dist.ReduceOp.SUM)
            return [float(tmax[0])] + [float(x) for x in tsum[1:]]
        return [float(x) for x in t]

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


LAYOUTS = {"reference": 0, "wide": 1, "wide4": 2}
BUILDERS = {"device": "device", "ploc": "ploc", "sah": "sah", "split": "split", "split_lbvh": "split_lbvh"}


def optimize_setting(args, path_tracing):
    """(iterations, fraction) of the reinsertion pass for this run, or None."""
    v = getattr(args, "optimize_bvh", "auto")
    if v == "off" or (v == "auto" and not path_tracing):
        return None
    if v == "auto":
        return 4, 0.1
    it, fr = v.split(",")
    return int(it), float(fr)


def make_renderer(d, scene, W, H, wl, args, **kw):
    """One context per rank. N > 1: the tree is built ONCE, on rank 0, and reaches the other ranks through the library's
    broadcast (pt_comm_broadcast_bvh); RNG tables are seeded per rank by the library (pt_comm_seed_rng)."""
    from gpupathtracer_b200 import Renderer
    from gpupathtracer_b200.distributed import init_comm
    how = BUILDERS.get(args.builder, "build")
    if d.world > 1 and d.rank != 0 and LAYOUTS[args.bvh] == 0:
        how = "none"
    r = Renderer().Initialize(W, H, scene, max_depth=wl["depth"], microfacet_model=wl["model"], device=d.local, bvh=how,
                              bvh_layout=LAYOUTS[args.bvh], **kw)
    opt = optimize_setting(args, True)
    if opt and how != "none" and LAYOUTS[args.bvh] == 0:
        r.OptimizeBVH(*opt)                       # on the rank that built the tree; the others receive the optimised tree below
    if d.world > 1:
        init_comm(r, d.dist)
        if LAYOUTS[args.bvh] == 0:
            r.CommBroadcastBVH(0)
            if d.rank != 0:
                scene.set_bvh(*r.GetBVH())
        r.CommSeedRNG(0xB200)
    else:
        r.lib.pt_seed_rng(r.ctx, 0xB200)
    return r


def rel_mse(a, b, eps=1e-2):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    ok = np.isfinite(a) & np.isfinite(b)
    return float(np.mean(((a - b) ** 2 / (b ** 2 + eps))[ok]))


def parity_sample(scene, wl, frames=2):
    """The CUDA path against the reference's own main() on a bounded sample of THIS workload (the 480 x 272 crop the CPU arm
    renders, same camera / scene / tree / depth / RNG table): bit-for-bit under the parity libm (both sides evaluate sin / cos /
    log / exp through csrc/pt_parity_libm.h), and how many pixels the default fast-libm build differs on."""
    import oracle
    from gpupathtracer_b200 import Renderer, make_rng_table
    if not oracle.have_glsl():
        return {"skipped": "oracle/_ref/libref_glsl.so not built"}
    w, h = CPU_SAMPLE
    cam = make_camera(scene, w, h, aspect=wl["width"] / wl["height"])
    rng0 = make_rng_table(w * h, 4321)
    o = oracle.Oracle(scene)
    out = {"pixels": w * h, "frames": frames, "against": "the reference's main() (Iterative.comp:214-307 compiled as C++, all host cores)"}
    for plibm, code in ((True, 6 if wl["model"] else 4), (False, 2 if wl["model"] else 0)):
        r = Renderer().Initialize(w, h, scene, max_depth=wl["depth"], microfacet_model=wl["model"], rng_state=rng0, bvh="given", parity_libm=plibm)
        r.RenderFrame(cam, frames)
        got, rng_got = r.DownloadAccum(), r.GetRNG()
        r.CleanUp()
        rng_ref, want = rng0.copy(), np.zeros((h, w, 4), np.float32)
        o.ref_glsl_render_mp(cam.as_array(), w, h, rng_ref, want, n_frames=frames, max_depth=wl["depth"], model=code)
        same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
        bad = int((~same.all(axis=2)).sum())
        if plibm:
            out.update({"accum_bit_exact": bad == 0, "rng_bit_exact": bool(np.array_equal(rng_got, rng_ref)), "mismatching_pixels": bad})
        else:
            close = np.isclose(got[..., :3], want[..., :3], rtol=1e-4, atol=1e-6).all(axis=2) | (np.isnan(got[..., :3]) & np.isnan(want[..., :3])).all(axis=2)
            out["default_libm"] = {"pixels_differing_in_some_bit": bad / float(w * h), "pixels_differing_by_more_than_1e-4": float((~close).mean())}
    return out


def timed_batches(d, r, cam, steps, min_seconds=1.0, max_reps=64):
    """Repeat the K-frame batch until the timed region is at least `min_seconds` long; every repetition is bracketed by a
    barrier + synchronize and timed on the device (CUDA events on the library's streams, max over ranks; for N > 1 the
    batch includes its reduce to rank 0). Returns the per-repetition records and the clock record."""
    reps, total_ms = [], 0.0
    sampler = ClockSampler(d.local); sampler.start()
    t_wall = time.perf_counter()
    while len(reps) < 3 or (total_ms < min_seconds * 1e3 and len(reps) < max_reps):
        d.barrier(r)
        c0 = r.Counters()
        if d.world > 1:
            r.RenderBatch(cam, steps); r.CommSynchronize()
            ms = r.CommLastBatchMs()
        else:
            r.RenderFrame(cam, steps); r.Synchronize()
            ms = r.LastDeviceMs()
        d.barrier(r)
        c1 = r.Counters()
        ms, closest, shadow, nulls = d.max_sum([ms, float(c1["closest_rays"] - c0["closest_rays"]), float(c1["shadow_rays"] - c0["shadow_rays"]),
                                                float(c1["null_shadow_rays"] - c0["null_shadow_rays"])])
        reps.append((ms, closest, shadow, nulls, int(c1["kernel_launches"] - c0["kernel_launches"])))
        total_ms += ms
    return reps, sampler.result(), time.perf_counter() - t_wall


def all_nee_pass(d, scene, W, H, wl, args, cam, steps, warmup):
    """The same K frames with every NEE sample traced, as the reference and the CPU arm do (pt_config.trace_null_shadow):
    identical image, the ray population of the reference; reported beside `value`, not instead of it."""
    from gpupathtracer_b200 import Renderer
    r_all = Renderer().Initialize(W, H, scene, max_depth=wl["depth"], microfacet_model=wl["model"], device=d.local, seed=0xB200,
                                  bvh="given", bvh_layout=LAYOUTS[args.bvh], trace_null_shadow=True)
    for _ in range(warmup):
        r_all.RenderFrame(cam, 1)
    r_all.Synchronize()
    best = None
    for _ in range(3):
        ca0 = r_all.Counters()
        r_all.RenderFrame(cam, steps); r_all.Synchronize()
        ms = r_all.LastDeviceMs()
        ca1 = r_all.Counters()
        rays = float(ca1["closest_rays"] - ca0["closest_rays"] + ca1["shadow_rays"] - ca0["shadow_rays"])
        if best is None or ms < best[0]:
            best = (ms, rays)
    r_all.CleanUp()
    ms, rays = best
    return {"value": rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "samples_per_s": steps * W * H / (ms * 1e-3), "ms_per_step": ms / steps,
            "note": "pt_config.trace_null_shadow = 1: NEE samples with exactly zero radiance are traced too, the reference's ray population"}


def run_pathtrace(args, name, wl, d, sub=False):
    """One path-tracing workload -> the JSON line as a dict. sub: a short run for the `configs` block of the default command
    (no all-NEE pass, no CPU baseline)."""
    torch, world, rank, local = d.torch, d.world, d.rank, d.local
    W, H, depth = wl["width"], wl["height"], wl["depth"]
    scene = build_scene(wl)
    t0 = time.perf_counter()
    r = make_renderer(d, scene, W, H, wl, args)
    init_s = time.perf_counter() - t0
    cam = make_camera(scene, W, H)
    pinned = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
    warmup = max(args.warmup, 3)
    steps = args.steps

    def render(n):
        r.RenderBatch(cam, n) if world > 1 else r.RenderFrame(cam, n)

    for _ in range(warmup):
        render(1)
    d.barrier(r)
    r.CommReset() if world > 1 else r.ResetSamples()
    # ---- timed region: repetitions of K frames back to back, device time; scene / BVH / RNG tables resident in HBM
    reps, clocks, region_s = timed_batches(d, r, cam, steps, min_seconds=0.4 if sub else 1.0)
    ms_all_reps = sorted(x[0] for x in reps)
    ms_total, closest, shadow, nulls, launches = sorted(reps, key=lambda x: x[0])[len(reps) // 2]
    rays = closest + shadow
    value = rays / (ms_total * 1e-3) / 1e6
    samples_per_s = world * steps * W * H / (ms_total * 1e-3)
    ms_reduce = r.CommLastReduceMs() if world > 1 else 0.0

    # ---- end to end through the public API with host buffers: camera in (H2D), the presented 8-bit frame out (D2H).
    #      N > 1: every step is a one-frame batch on every rank + its reduce, and rank 0 presents the COMBINED image.
    d.barrier(r)
    c2 = r.Counters()
    e2e_steps = steps if not sub else min(steps, 16)
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        cam.GenerateImagePlane()                   # host: Camera::GenerateImagePlane; H2D: the 76 B of camera uniforms
        if world > 1:
            r.RenderBatch(cam, 1)
            r.CommPresentAsync(pinned[i & 1].data_ptr())
        else:
            r.RenderFrame(cam, 1)
            r.PresentAsync(pinned[i & 1].data_ptr())   # Present(): /numSamples + tonemap, D2H into pinned host memory (double-buffered)
    d.barrier(r)                                   # all K presented frames are in host memory here
    e2e_s = time.perf_counter() - t0
    c3 = r.Counters()
    e2e_s, e2e_rays = d.max_sum([e2e_s, float(c3["closest_rays"] - c2["closest_rays"] + c3["shadow_rays"] - c2["shadow_rays"])])
    e2e_value = e2e_rays / e2e_s / 1e6

    # ---- per-kernel durations: a separate pass with the frame as ONE launch sequence (no overlapping slices), CUDA events
    #      between kernels; then two frames through the instrumented kernels for the visit counts
    stage_frames = min(steps, 8)
    c4 = r.Counters()
    r.SetStageTiming(True)
    r.RenderFrame(cam, stage_frames); r.Synchronize()
    stage = r.StageMs()
    r.SetStageTiming(False)
    c5 = r.Counters()
    closest_stage = float(c5["closest_rays"] - c4["closest_rays"])
    shadow_stage = float(c5["shadow_rays"] - c4["shadow_rays"])
    visits = None
    if args.bvh == "reference":
        r.SetVisitCounting(True)
        r.RenderFrame(cam, 2); r.Synchronize()
        r.SetVisitCounting(False)
        v = r.Counters()
        visits = {"closest_pairs_per_ray": v["closest_pair_tests"] / max(v["counted_closest_rays"], 1),
                  "closest_tris_per_ray": v["closest_tri_tests"] / max(v["counted_closest_rays"], 1),
                  "shadow_pairs_per_ray": v["shadow_pair_tests"] / max(v["counted_shadow_rays"], 1),
                  "shadow_tris_per_ray": v["shadow_tri_tests"] / max(v["counted_shadow_rays"], 1),
                  "source": "device counters of the instrumented traversal kernels (pt_set_visit_counting) over 2 frames of this workload"}
    line = None
    if rank == 0:
        line = pathtrace_line(args, name, wl, d, scene, r, cam, dict(
            value=value, samples_per_s=samples_per_s, closest=closest, shadow=shadow, nulls=nulls, launches=launches, steps=steps,
            warmup=warmup, ms_total=ms_total, ms_all_reps=ms_all_reps, reps=len(reps), region_s=region_s, clocks=clocks, ms_reduce=ms_reduce,
            e2e_value=e2e_value, e2e_s=e2e_s, e2e_steps=e2e_steps, stage=stage, stage_frames=stage_frames, closest_stage=closest_stage,
            shadow_stage=shadow_stage, visits=visits, init_s=init_s))
        if world == 1 and not sub:
            line["all_nee_samples_traced"] = all_nee_pass(d, scene, W, H, wl, args, cam, steps, warmup)
        if world == 1 and not args.no_cpu_baseline:
            line["parity_on_sample"] = parity_sample(scene, wl)
        if world == 1 and not args.no_cpu_baseline and not sub:
            w, h = CPU_SAMPLE
            mrays, sps, ms, tot, info = cpu_reference_run(scene, 4, 1, w, h, wl)
            line["cpu_baseline"] = {"value": mrays, "unit": "Mrays/s", "cores": os.cpu_count() or 1, "kind": info["kind"],
                                    "sample": f"{w}x{h} px of the same camera/scene/tree, 4 frames; {info['what']}",
                                    "samples_per_s": sps, **{k: v for k, v in info.items() if k not in ("kind", "what")}}
    r.CleanUp()
    return line


def pathtrace_line(args, name, wl, d, scene, r, cam, m):
    """The JSON line of a path-tracing workload from the measurements `m` of run_pathtrace."""
    world = d.world
    W, H, depth, steps = wl["width"], wl["height"], wl["depth"], m["steps"]
    peak, which = measured_peaks()
    line = {
        "metric": metric_name(name, wl), "value": m["value"], "unit": "Mrays/s", "samples_per_s": m["samples_per_s"],
        "rays_per_step": {"closest": m["closest"] / steps, "shadow_traced": m["shadow"] / steps, "shadow_not_needed": m["nulls"] / steps,
                          "note": "NEE samples whose radiance is exactly 0 (light below the horizon) are settled without a ray -- the "
                                  "reference traces them and adds 0; value counts traced rays only"},
        "n_gpus": world, "steps": steps, "warmup": m["warmup"], "ms_per_step": m["ms_total"] / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "timed_region": {"repetitions": m["reps"], "steps_each": steps, "seconds": sum(m["ms_all_reps"]) * 1e-3, "wall_s": m["region_s"],
                         "ms_per_step_min": m["ms_all_reps"][0] / steps, "ms_per_step_median": m["ms_total"] / steps,
                         "ms_per_step_max": m["ms_all_reps"][-1] / steps,
                         "note": "value / ms_per_step are the MEDIAN repetition of K back-to-back frames"},
        "config": {"workload": wl["desc"], "width": W, "height": H, "max_depth": depth, "triangles": scene.tris.shape[0],
                   "microfacet": "beckmann" if wl["model"] else "ggx", "lens_radius": float(cam.plane.lens_radius),
                   "bvh": f"{ {'device': 'device LBVH', 'ploc': 'device PLOC', 'sah': 'device top-down SAH', 'split': 'device pre-split + PLOC', 'split_lbvh': 'device pre-split + LBVH'}.get(args.builder, 'own SBVH') } builder, {args.bvh} layout",
                   "bvh_total_ms": float(r.bvh_stats.total_ms), "bvh_sah_cost": float(r.bvh_stats.sah_cost), "bvh_build_ms": float(r.bvh_stats.build_ms),
                   "bvh_optimized": ({"what": "pt_optimize_bvh: reinsertion pass over the built tree (host), same leaves, same hits",
                                      "iterations_fraction": list(optimize_setting(args, True)), "sah_cost": float(r.opt_stats.sah_cost),
                                      "ms": float(r.opt_stats.total_ms), "depth": int(r.opt_stats.max_depth)}
                                     if getattr(r, "opt_stats", None) is not None else None),
                   "slices": "each frame = 3 pixel-disjoint slices on concurrent streams, shadow(b) || extend(b+1) inside a slice (pt_config.n_slices)",
                   "l2": "per-step working set (path state + queues + accum, ~0.5 GB at 1080p) exceeds the 126 MB L2",
                   "parallelism": (f"samples split over {world} GPUs inside the library (pt_comm_init): BVH built on rank 0 and broadcast, "
                                   f"per-rank RNG seeds, 1 ncclReduce of the accumulation buffer per {steps}-frame batch on its own stream"
                                   if world > 1 else "1 GPU"),
                   "init_s": m["init_s"]},
        "e2e": {"value": m["e2e_value"], "unit": "Mrays/s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": W * H * 3,
                "ms_per_step": m["e2e_s"] / m["e2e_steps"] * 1e3, "steps": m["e2e_steps"],
                "what": ("RenderFrame + Present per step: camera in, every rank renders a frame, ncclReduce to rank 0, the COMBINED "
                         "frame resolved and copied to rank 0's pinned host memory" if world > 1 else
                         "RenderFrame + PresentAsync per step through the C ABI: camera in, tonemapped 8-bit frame out to pinned host "
                         "memory; slices chain frame to frame (no per-frame drain), each slice resolves its own pixels")},
        "gpu_launches": m["launches"], "clocks": m["clocks"], "ms_reduce": m["ms_reduce"],
        "ms_reduce_note": "first band's ncclReduce to the last band's on the communication stream: includes waiting for the later slices (chained batches)" if world > 1 else None,
    }
    stage, stage_frames, visits = m["stage"], m["stage_frames"], m["visits"]
    line["stage_ms_per_step"] = {k: v / stage_frames for k, v in stage.items()}
    line["stage_note"] = (f"per-kernel times from a separate {stage_frames}-frame pass without slice overlap; their sum exceeds "
                          "ms_per_step because the timed region overlaps slices")
    ext_s, n_launch = stage["extend"] * 1e-3, stage_frames * depth
    if visits and ext_s > 0:
        pairs_c, tris_c = visits["closest_pairs_per_ray"], visits["closest_tris_per_ray"]
        b_ray = B_RAY_IN + B_HIT_OUT + pairs_c * B_PAIR + tris_c * B_TRI
        achieved = m["closest_stage"] * b_ray / ext_s / 1e9
        compulsory = m["closest_stage"] * (B_RAY_IN + B_HIT_OUT) / ext_s / 1e9
        tr = traffic_for(name, "k_extend")
        dram_per_launch = tr.get("dram_bytes_per_launch") if isinstance(tr, dict) else tr
        line["roofline"] = {
            "bound": "hbm", "kernel": "k_extend", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": dram_per_launch, "compulsory_frac": compulsory / peak,
            "dram_frac": (dram_per_launch / (stage["extend"] / n_launch * 1e-3) / 1e9 / peak) if dram_per_launch else None,
            "binding_limit": "L1 data pipe / SIMD efficiency, not HBM: the ~50 MB scene is L2-resident, so `frac` (algorithmic bytes over the "
                             "HBM peak, as the contract defines it) can exceed 1; compulsory_frac counts ray-in + hit-out only, dram_frac the "
                             "DRAM bytes ncu measured (profiles/traffic.json)",
            "ncu": traffic_for(name, "k_extend ncu"),
            "algorithmic_bytes_per_launch": m["closest_stage"] * b_ray / n_launch, "peak_source": which,
            "bytes_per_ray": b_ray, "pairs_per_ray": pairs_c, "tris_per_ray": tris_c, "visit_counts": visits,
            "launches": n_launch, "avg_launch_ms": stage["extend"] / n_launch,
            "shadow": {"avg_launch_ms": stage["shadow"] / n_launch, "rays_per_launch": m["shadow_stage"] / n_launch,
                       "bytes_per_ray": B_RAY_IN + B_FLAG_OUT + visits["shadow_pairs_per_ray"] * B_PAIR + visits["shadow_tris_per_ray"] * B_TRI}}
    return line


def run_fixed_spp(args, name, wl, d):
    """BASELINE.json configs[4] as written: a FIXED total of samples per pixel (--spp, 4096) split over the GPUs -- spp / N frames
    per GPU in 64-frame batches, one ncclReduce per batch overlapped with the next batch -- and the time until the combined,
    tonemapped image is in rank 0's host memory. The reduced image is then checked against two further, independent batches."""
    torch, world, rank, local = d.torch, d.world, d.rank, d.local
    W, H, depth = wl["width"], wl["height"], wl["depth"]
    per_gpu = (args.spp + world - 1) // world
    total, batch = per_gpu * world, 64
    scene = build_scene(wl)
    r = make_renderer(d, scene, W, H, wl, args)
    cam = make_camera(scene, W, H)
    pinned = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()

    def render(n):
        r.RenderBatch(cam, n) if world > 1 else r.RenderFrame(cam, n)

    def total_image():
        if world > 1:
            r.CommSynchronize()
            return r.CommDownloadTotal()[..., :3] if rank == 0 else None
        r.Synchronize()
        return r.DownloadAccum()[..., :3]

    for _ in range(max(args.warmup, 3)):
        render(1)
    d.barrier(r)
    r.CommReset() if world > 1 else r.ResetSamples()
    d.barrier(r)
    c0 = r.Counters()
    sampler = ClockSampler(local); sampler.start()
    t0 = time.perf_counter()
    done, n_batches = 0, 0
    while done < per_gpu:
        n = min(batch, per_gpu - done)
        render(n)
        done += n; n_batches += 1
    r.CommPresentAsync(pinned.data_ptr()) if world > 1 else r.PresentAsync(pinned.data_ptr())
    d.barrier(r)                                   # the combined 8-bit image is in rank 0's host memory
    wall = time.perf_counter() - t0
    clocks = sampler.result()
    c1 = r.Counters()
    wall, rays = d.max_sum([wall, float(c1["closest_rays"] - c0["closest_rays"] + c1["shadow_rays"] - c0["shadow_rays"])])
    final = total_image()
    spp_final = r.CommNumSamples() if world > 1 else r.GetNumSamples()
    render(batch); t1 = total_image()              # image check (untimed): two more batches, each an independent estimate
    render(batch); t2 = total_image()
    line = None
    if rank == 0:
        nb = batch * world
        a_img, b_img, f_img = (t1 - final) / nb, (t2 - t1) / nb, final / spp_final
        floor = rel_mse(a_img, b_img)
        err = 0.5 * (rel_mse(a_img, f_img) + rel_mse(b_img, f_img))
        line = {"metric": f"time to image, {name} {W}x{H} depth {depth}, {total} spp split over the GPUs, reduce per {batch}-frame batch",
                "value": wall, "unit": "s", "higher_is_better": False, "n_gpus": world, "steps": n_batches, "warmup": max(args.warmup, 3),
                "ms_per_step": wall / n_batches * 1e3, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "samples_per_s": total * W * H / wall, "mrays_per_s": rays / wall / 1e6,
                "config": {"workload": wl["desc"], "width": W, "height": H, "total_spp": total, "frames_per_gpu": per_gpu, "batches": n_batches,
                           "triangles": scene.tris.shape[0], "parallelism": f"{world} GPU(s), samples split inside the library, one ncclReduce per "
                           "batch on its own stream overlapped with the next batch (double-buffered accumulators)"},
                "e2e": {"value": wall, "unit": "s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": W * H * 3 / n_batches,
                        "what": "wall clock from the first batch to the combined tonemapped image in rank 0's pinned host memory"},
                "image_check": {"spp": int(spp_final), "relmse_final_vs_independent_batches": err, "relmse_between_independent_batches": floor,
                                "ok": bool(err < floor and np.isfinite(f_img).mean() > 0.999),
                                "note": f"two further {nb}-spp batches (untimed) are each an independent estimate of the image; the final image "
                                        "must be closer to both than they are to each other"},
                "reduce_ms_last": r.CommLastReduceMs() if world > 1 else 0.0,
                "gpu_launches": int(c1["kernel_launches"] - c0["kernel_launches"]), "clocks": clocks}
    r.CleanUp()
    return line


def run_micro(args, name, wl, d, sub=False):
    """configs[3]: fixed ray batches against a scene far larger than L2 -- the HBM-bound case."""
    torch, world, rank, local = d.torch, d.world, d.rank, d.local
    line = None
    from gpupathtracer_b200 import Renderer
    n = wl["n_rays"]
    scene, closest, anyr = micro_scene_and_rays(n, 0xB200 + 7 * rank)
    t0 = time.perf_counter()
    r = Renderer().Initialize(64, 64, scene, max_depth=5, device=local, seed=1, bvh=BUILDERS.get(args.builder, "build"),
                              bvh_layout=LAYOUTS[args.bvh])
    if optimize_setting(args, False) and LAYOUTS[args.bvh] == 0:
        r.OptimizeBVH(*optimize_setting(args, False))
    init_s = time.perf_counter() - t0
    d_closest = torch.from_numpy(closest).cuda()
    d_any = torch.from_numpy(anyr).cuda()
    d_hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    d_flags = torch.empty((n,), dtype=torch.uint8, device="cuda")
    h_closest = torch.from_numpy(closest).pin_memory()
    h_any = torch.from_numpy(anyr).pin_memory()
    h_hits = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    h_flags = torch.empty((n,), dtype=torch.uint8).pin_memory()
    torch.cuda.synchronize()
    warmup = max(args.warmup, 3)
    steps = args.steps

    def step_device():
        r.TraceClosestDevice(d_closest.data_ptr(), n, d_hits.data_ptr()); ms_c = r.LastDeviceMs()
        r.TraceAnyDevice(d_any.data_ptr(), n, d_flags.data_ptr()); ms_a = r.LastDeviceMs()
        return ms_c, ms_a

    for _ in range(warmup):
        step_device()
    d.barrier(r)
    c0 = r.Counters()
    sampler = ClockSampler(local); sampler.start()
    ms_c = ms_a = 0.0
    for _ in range(steps):                         # rays (2 x 512 MB) and tree (1.2 GB) are far larger than L2
        a, b = step_device(); ms_c += a; ms_a += b
    d.barrier(r)
    clocks = sampler.result()
    c1 = r.Counters()
    ms_total, rays = d.max_sum([ms_c + ms_a, 2.0 * n * steps])
    value = rays / (ms_total * 1e-3) / 1e6

    # ---- end to end: host ray batches in (pinned), hits / flags out, through pt_trace_closest / pt_trace_any
    d.barrier(r)
    t0 = time.perf_counter()
    e2e_steps = min(steps, 4)
    for _ in range(e2e_steps):
        r.TraceClosestHost(h_closest.data_ptr(), n, h_hits.data_ptr())
        r.TraceAnyHost(h_any.data_ptr(), n, h_flags.data_ptr())
    d.barrier(r)
    e2e_s, e2e_rays = d.max_sum([time.perf_counter() - t0, 2.0 * n * e2e_steps])

    if rank == 0:
        peak, which = measured_peaks()
        hits = d_hits.cpu().numpy()
        line = {
            "metric": metric_name(name, wl), "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_total / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl["desc"], "triangles": scene.tris.shape[0], "rays_per_step": 2 * n,
                       "bvh": f"{ {'device': 'device LBVH', 'ploc': 'device PLOC', 'sah': 'device top-down SAH', 'split': 'device pre-split + PLOC', 'split_lbvh': 'device pre-split + LBVH'}.get(args.builder, 'own SBVH') } builder, {args.bvh} layout",
                       "bvh_total_ms": float(r.bvh_stats.total_ms), "bvh_sah_cost": float(r.bvh_stats.sah_cost), "bvh_build_ms": float(r.bvh_stats.build_ms),
                       "bvh_nodes": int(r.bvh_stats.n_nodes), "bvh_refs": int(r.bvh_stats.n_refs),
                       "l2": "ray batches (2 x 512 MB) and tree + leaf stream (~1.2 GB) exceed the 126 MB L2",
                       "parallelism": f"{world} independent replica(s), each with its own 16M-ray batch; no collective", "init_s": init_s},
            "closest_mrays_s": n * steps / (ms_c * 1e-3) / 1e6, "any_mrays_s": n * steps / (ms_a * 1e-3) / 1e6,
            "hit_fraction": float((hits[:, 3].view(np.int32) >= 0).mean()),
            "e2e": {"value": e2e_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 2 * n * 32, "d2h_bytes_per_step": n * 16 + n,
                    "ms_per_step": e2e_s / e2e_steps * 1e3},
            "gpu_launches": int(c1["kernel_launches"] - c0["kernel_launches"]),
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            import oracle
            o = oracle.Oracle(scene)
            m = 1 << 19
            t0 = time.perf_counter()
            want, cc = o.trace_closest(closest[:m])
            wflag, ca = o.trace_any(anyr[:m])
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": 2 * m / dt / 1e6, "unit": "Mrays/s", "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": f"first {m} closest-hit + {m} any-hit rays of the same batch, all host cores"}
            line["parity_on_sample"] = {"closest_bit_exact": bool(np.array_equal(want.view(np.uint32), hits[:m].view(np.uint32))),
                                        "any_bit_exact": bool(np.array_equal(wflag, d_flags[:m].cpu().numpy()))}
            b_ray, pairs, tris = bytes_per_ray(cc, m, B_HIT_OUT)
            if args.bvh == "reference":                      # visit counts of ALL rays of the batch, from the instrumented kernel
                r.SetVisitCounting(True)
                r.TraceClosestDevice(d_closest.data_ptr(), n, d_hits.data_ptr()); r.Synchronize()
                r.SetVisitCounting(False)
                v = r.Counters()
                pairs, tris = v["closest_pair_tests"] / v["counted_closest_rays"], v["closest_tri_tests"] / v["counted_closest_rays"]
                b_ray = B_RAY_IN + B_HIT_OUT + pairs * B_PAIR + tris * B_TRI
            achieved = n * b_ray / (ms_c / steps * 1e-3) / 1e9
            line["roofline"] = {"bound": "hbm", "kernel": "k_trace_batch<false>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                                "frac": achieved / peak,
                                "traffic": traffic_for(name, "k_trace_batch<false>" + {"wide": " wide-8", "wide4": " wide-4"}.get(args.bvh, "")),
                                "ncu": traffic_for(name, "k_trace_batch<false> ncu") if args.bvh == "reference" else None,
                                "algorithmic_bytes_per_launch": n * b_ray, "peak_source": which, "bytes_per_ray": b_ray,
                                "pairs_per_ray": pairs, "tris_per_ray": tris, "launches": steps, "avg_launch_ms": ms_c / steps,
                                "note": "avg_launch_ms spans the ray binning (4 small kernels, ~0.35 ms, reference layout only) and "
                                        "k_trace_batch; the binning halves the DRAM traffic of the trace (profiles/traffic.json)"}
    r.CleanUp()
    return line


def run_primary(args, name, wl, d, sub=False):
    """configs[0]: BVH build + 1 spp primary closest-hit at 1280x720."""
    torch, world, rank, local = d.torch, d.world, d.rank, d.local
    line = None
    from gpupathtracer_b200 import Renderer, Scene, scenes, build_bvh_host
    W, H = wl["width"], wl["height"]
    scene = Scene(scenes.bunny_like(seed=1))
    builds = []
    for _ in range(3):
        _, _, st = build_bvh_host(scene.tris)
        builds.append(st["build_ms"])
    r = Renderer().Initialize(W, H, scene, max_depth=1, device=local, seed=0xB200 + rank,
                              bvh=BUILDERS.get(args.builder, "build"), bvh_layout=LAYOUTS[args.bvh])
    cam = make_camera(scene, W, H)
    r.lib.pt_set_camera(r.ctx, cam.plane)
    n = W * H
    # the device builders (Morton LBVH, PLOC, top-down binned SAH): build time and what their trees cost to traverse
    dev = {}
    if args.bvh == "reference":
        import ctypes as C
        from gpupathtracer_b200.api import pt_bvh_stats
        for key, which in (("host_sbvh", "build"), ("device_lbvh", "device"), ("device_ploc", "ploc"), ("device_sah", "sah")):
            s2 = Scene(scene.data)
            r2 = Renderer().Initialize(W, H, s2, max_depth=1, device=local, seed=1, bvh=which)
            dev_builds = [float(r2.bvh_stats.build_ms)]
            tot = float(r2.bvh_stats.total_ms)
            for _ in range(4 if which != "build" else 0):
                st = pt_bvh_stats()
                if which == "ploc":
                    r2.lib.pt_build_bvh_device_ploc(r2.ctx, 16, C.byref(st))
                elif which == "sah":
                    r2.lib.pt_build_bvh_device_sah(r2.ctx, C.byref(st))
                else:
                    r2.lib.pt_build_bvh_device(r2.ctx, C.byref(st))
                dev_builds.append(float(st.build_ms)); tot = float(st.total_ms)
            dev.update({key + "_kernels_best_of_5": min(dev_builds), key + "_total_call_ms": tot,
                        key + "_sah_cost": float(r2.bvh_stats.sah_cost), key + "_depth": int(r2.bvh_stats.max_depth)})
            r2.lib.pt_set_camera(r2.ctx, cam.plane)
            pr2 = r2.RaygenProbe(0, n)
            rays2 = np.zeros((n, 8), np.float32); rays2[:, 0:3] = pr2[:, 2:5]; rays2[:, 4:7] = pr2[:, 5:8]; rays2[:, 7] = 1e20
            d2 = torch.from_numpy(rays2).cuda(); h2 = torch.empty((n, 4), dtype=torch.float32, device="cuda"); torch.cuda.synchronize()
            ms2 = []
            for _ in range(8):
                r2.TraceClosestDevice(d2.data_ptr(), n, h2.data_ptr()); ms2.append(r2.LastDeviceMs())
            dev[key + "_primary_mrays_s_warm_l2"] = n / min(ms2) / 1e3     # best of 8 back-to-back launches: the tree stays in L2
            fl = torch.empty(256 << 20, dtype=torch.uint8, device="cuda"); ms3 = []
            for _ in range(8):
                fl.zero_(); torch.cuda.synchronize()
                r2.TraceClosestDevice(d2.data_ptr(), n, h2.data_ptr()); ms3.append(r2.LastDeviceMs())
            dev[key + "_primary_mrays_s_flushed_l2"] = n * len(ms3) / sum(ms3) / 1e3    # as the timed region does
            del fl
            r2.CleanUp()
    pr = r.RaygenProbe(0, n)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = pr[:, 2:5]; rays[:, 4:7] = pr[:, 5:8]; rays[:, 7] = 1e20
    d_rays = torch.from_numpy(rays).cuda()
    d_hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    h_rays = torch.from_numpy(rays).pin_memory(); h_hits = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    torch.cuda.synchronize()
    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        r.TraceClosestDevice(d_rays.data_ptr(), n, d_hits.data_ptr())
    d.barrier(r)
    sampler = ClockSampler(local); sampler.start()
    ms = 0.0
    c0 = r.Counters()
    for _ in range(args.steps):
        flush.zero_(); torch.cuda.synchronize()    # 256 MB write between timed launches: L2 flushed
        r.TraceClosestDevice(d_rays.data_ptr(), n, d_hits.data_ptr()); ms += r.LastDeviceMs()
    d.barrier(r)
    clocks = sampler.result()
    c1 = r.Counters()
    ms_total, rays_total = d.max_sum([ms, float(n * args.steps)])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r.TraceClosestHost(h_rays.data_ptr(), n, h_hits.data_ptr())
    d.barrier(r)
    e2e_s, e2e_rays = d.max_sum([time.perf_counter() - t0, float(n * args.steps)])
    # the same rays generated on the device from the camera (pt_trace_primary_async = InitRay + closest hit of RenderFrame's
    # first bounce): camera in, hit records in pixel order out, two pinned host buffers, transfer overlapped with the next trace
    h_prim = [torch.empty((n, 4), dtype=torch.float32).pin_memory() for _ in range(2)]
    for k in range(2):
        r.TracePrimaryAsync(h_prim[k].data_ptr())
    r.Synchronize()
    d.barrier(r)
    t0 = time.perf_counter()
    for k in range(args.steps):
        r.lib.pt_set_camera(r.ctx, cam.plane)
        r.TracePrimaryAsync(h_prim[k & 1].data_ptr())
    r.Synchronize()
    d.barrier(r)
    cam_s, cam_rays = d.max_sum([time.perf_counter() - t0, float(n * args.steps)])
    if rank == 0:
        peak, which = measured_peaks()
        hits = d_hits.cpu().numpy()
        line = {
            "metric": metric_name(name, wl), "value": rays_total / (ms_total * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "triangles": scene.tris.shape[0], "width": W, "height": H,
                       "bvh": {"device": "device LBVH builder", "ploc": "device PLOC builder (pt_build_bvh_device_ploc, radius 16)",
                               "sah": "device top-down binned SAH builder (pt_build_bvh_device_sah)"}.get(
                           args.builder, "own SBVH builder (host, all cores)") + f", {args.bvh} layout",
                       "bvh_build_kernels_ms": float(r.bvh_stats.build_ms) if args.builder in ("device", "ploc", "sah") else None,
                       "l2": "256 MB flush between timed launches"},
            "bvh_build_ms": {**dev, "ours_host_all_cores_best_of_3": min(builds), "nodes": int(r.bvh_stats.n_nodes), "refs": int(r.bvh_stats.n_refs),
                             "sah_cost": float(r.bvh_stats.sah_cost)},
            "hit_fraction": float((hits[:, 3].view(np.int32) >= 0).mean()),
            "e2e": {"value": cam_rays / cam_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": n * 16,
                    "ms_per_step": cam_s / args.steps * 1e3,
                    "note": "pt_trace_primary_async per step: camera in (76 B), rays generated on the device as RenderFrame's first bounce "
                            "does (InitRay + closest hit; the reference has no host ray batches), hit records in pixel order into pinned "
                            "host memory, transfer overlapped with the next trace; all frames in host memory at the closing barrier"},
            "e2e_host_ray_batches": {"value": e2e_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": n * 16,
                                     "ms_per_step": e2e_s / args.steps * 1e3, "note": "pt_trace_closest on a host batch: 32 B per ray over PCIe first"},
            "gpu_launches": int(c1["kernel_launches"] - c0["kernel_launches"]),
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            import oracle
            o = oracle.Oracle(scene)
            t0 = time.perf_counter()
            want, cc = o.trace_closest(rays)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt / 1e6, "unit": "Mrays/s", "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": "all 921 600 primary rays, same tree, all host cores"}
            line["parity_on_sample"] = {"closest_bit_exact": bool(np.array_equal(want.view(np.uint32), hits.view(np.uint32)))}
            if oracle.have_ref():
                t0 = time.perf_counter()
                oracle.ref_build_sbvh(scene.tris_abs)
                line["bvh_build_ms"]["reference_BuildBinnedSpatial_1_thread"] = (time.perf_counter() - t0) * 1e3
            b_ray, pairs, tris = bytes_per_ray(cc, n, B_HIT_OUT)
            achieved = n * b_ray / (ms_total / args.steps * 1e-3) / 1e9
            line["roofline"] = {"bound": "hbm", "kernel": "k_trace_batch<false>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                                "frac": achieved / peak, "traffic": traffic_for(name, "k_trace_batch<false>"),
                                "ncu": traffic_for(name, "k_trace_batch<false> ncu"),
                                "algorithmic_bytes_per_launch": n * b_ray, "peak_source": which, "bytes_per_ray": b_ray,
                                "pairs_per_ray": pairs, "tris_per_ray": tris, "launches": args.steps, "avg_launch_ms": ms_total / args.steps,
                                "note": "coherent primary rays on a 10 MB tree: L1/L2-resident, issue-slot bound (DESIGN.md 4)"}
    r.CleanUp()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="atrium", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="default command only: skip the short runs of the other BASELINE.json configs")
    ap.add_argument("--spp", type=int, default=0, help="path-tracing workloads: render a FIXED total of samples per pixel split over the GPUs "
                    "(64-frame batches per GPU, one reduce per batch) and report time-to-image (BASELINE.json configs[4] as written: --workload 4k --spp 4096)")
    ap.add_argument("--optimize-bvh", default="auto", metavar="ITERATIONS,FRACTION",
                    help="reinsertion optimisation of the built tree (pt_optimize_bvh) before rendering: 'auto' = 4,0.1 for the path-tracing "
                         "workloads (measured +2-3 %% frames), off for primary-ray / random-ray batches (measured none); 'off'; or e.g. 8,0.3")
    ap.add_argument("--builder", default="host", choices=["host", "device", "ploc", "sah", "split", "split_lbvh"],
                    help="host: own multi-threaded SBVH builder (default); device: Morton LBVH built by CUDA kernels; ploc: the same with a "
                         "PLOC hierarchy (pt_build_bvh_device_ploc); sah: top-down binned SAH on the device (pt_build_bvh_device_sah)")
    ap.add_argument("--bvh", default="reference", choices=["reference", "wide", "wide4"],
                    help="device BVH layout: the reference's binary fp32 tree (bit-exact) or the compressed 8-/4-wide collapse of it")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    default_command = args.workload == "atrium" and args.builder == "host" and args.bvh == "reference" and not args.spp
    if args.steps is None:
        args.steps = {"pt": 64, "rays": 8, "primary": 32}[wl["kind"]] if args.impl != "reference" else 3
        if args.workload == "4k":
            args.steps = 16
    if args.impl == "reference":
        return run_reference(args)
    d = Dist()
    runners = {"pt": run_pathtrace, "rays": run_micro, "primary": run_primary}
    if args.spp:
        line = run_fixed_spp(args, args.workload, wl, d)
    else:
        line = runners[wl["kind"]](args, args.workload, wl, d)
    if default_command and d.world == 1 and not args.no_configs and line is not None:
        # the other BASELINE.json configs, short runs (value, e2e, parity_on_sample, roofline), so that one command observes all five
        import copy
        line["configs"] = {}
        for name, steps in (("bunny", 8), ("bathroom", 8), ("4k", 4), ("micro", 2)):
            a2 = copy.copy(args); a2.steps = steps; a2.warmup = 3
            t0 = time.perf_counter()
            try:
                sub = runners[WORKLOADS[name]["kind"]](a2, name, WORKLOADS[name], d, sub=True)
                sub["wall_s"] = time.perf_counter() - t0
                line["configs"][name] = sub
            except Exception as e:                 # never lose the headline line to a side run
                line["configs"][name] = {"error": repr(e)}
    if line is not None:
        print(json.dumps(line))
    d.close()


if __name__ == "__main__":
    main()
