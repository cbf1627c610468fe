#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on its config: Mrays/s (closest-hit + shadow rays actually
traced) and Msamples/s for "Sponza 1920x1080, 16 spp, 5 bounces" on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W                    # our arm
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle/_ref)

A step = one pass of the hot path over one frame: the whole 1920x1080 frame at 16 spp with 5 bounces
(genPath -> [closest-hit, shade, shadow-ray] x 5 -> gather; the 16 samples travel as one wavefront of 33 M paths).
Scene: the `sponza_atrium` blob (asset/sponza/sponza_lod.obj + sponza_lod.sbvh, Lambert/GGX, area
light + synthetic-sky IBL; the full-resolution Sponza OBJ is stripped from the reference snapshot,
SURVEY.md F10).  Synthetic data: seeds = std::mt19937(0), camera fixed.

value  : device-resident throughput; scene, seeds and path state already in HBM, CUDA events on the
         renderer's stream around the K steps, no host synchronisation inside (render calls are asynchronous).
e2e    : same metric through the public API with HOST buffers: per step the per-pixel seeds are
         copied host->device from pinned memory and the float4 image is read back to pinned memory.
N > 1  : one process per GPU (torchrun), scene replicated.  Default = STRONG scaling of the fixed frame: the 16
         samples of every pixel are split by sample index (rank r renders 16/N of them), the float4 sum images are
         combined with one NCCL reduce to rank 0 on the renderer's stream inside the timed region.
         --scaling weak: every rank renders its own 16 samples (a 16 N spp frame).  --partition tiles: interleaved
         64x64 tiles instead of sample ranges (same reduce; every pixel has one non-zero term).
roofline: the dominant kernels (the two traversal kernels, ~70 % of a step) are bound by instruction issue on an
         L2-resident scene, so the lead figure is the share of the SMs' lane-instruction rate that does useful work
         (issue-slot utilisation x active threads per instruction / 32, from the committed ncu summary of this
         build); the byte figures SURVEY 8(d) prescribes -- algorithmic bytes on the REFERENCE layout against the
         HBM copy peak, and the bytes that actually reach the L2 against a measured L2 read bandwidth -- follow.
"""
from __future__ import annotations

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

W, H, SPP, BOUNCE, RR = 1920, 1080, 16, 5, 3
SCENE = "sponza_atrium"
WORKLOAD = "sponza_lod 1920x1080 16spp 5-bounce NEE (configs[1] stand-in: full Sponza OBJ missing from snapshot)"
NCU_SUMMARY = os.path.join(ROOT, "profiles", "r02_ncu_summary.json")     # written by scripts/ncu_summary.py from a --set full capture


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# the reference's own CPU path
# ---------------------------------------------------------------------------------------------------------------
REF_W, REF_H, REF_SPP = 1920, 1080, 1      # the bench frame at 1 spp: only the sample count differs from the GPU arm's step


def reference_steps(budget_s: float, want: int, run):
    """Runs `run()` (one reference step, returns seconds and an info dict) at most `want` times within `budget_s`."""
    times, info = [], None
    t_all = time.perf_counter()
    while len(times) < want:
        s, info = run()
        times.append(s)
        if time.perf_counter() - t_all + s > budget_s:
            break
    return times, info


def reference_arm(args, rank: int, world: int):
    """The reference's own CPU path (aten::PathTracing over AcceleratedScene<sbvh>, compiled in place into
    oracle/_ref) on the box's host cores, all of them.  A step is a BOUNDED sample of the workload: the bench frame
    (same scene, camera, seeds, 1920x1080, 5 bounces) at 1 sample per pixel instead of 16 -- a rate, so comparable."""
    if rank != 0:
        return
    from tests import oracle_lib as ol
    cores = os.cpu_count() or 1
    if ol.have_ref():
        kind = "reference"
        # one harness process per step (scene build excluded: the harness times render() alone, like host_renderer does)
        def run():
            info = ol.ref_run("bench", SCENE, REF_W, REF_H, REF_SPP, BOUNCE, RR, cores, 1, timeout=3000)
            return info["best_ms"] / 1e3, info
        for _ in range(min(args.warmup, 1)):
            run()
        times, info = reference_steps(150.0, max(1, args.steps), run)
        ms = 1e3 * float(np.min(times))
        rays = info["mrays_per_s"] * info["best_ms"] * 1e3
        mrays = rays / (ms * 1e3)
        msamples = REF_W * REF_H * REF_SPP / (ms * 1e3)
    else:
        kind = "port"       # oracle/_ref did not travel: the restatement stands in and says so
        from aten_b200 import scene_io as sio
        scene = sio.load_scene(os.path.join(ROOT, "tests", "golden", f"{SCENE}.atnscene.gz")).with_size(REF_W, REF_H)
        state = {}
        def run():
            t0 = time.perf_counter()
            _, _, st = ol.render(scene, frames=REF_SPP, max_depth=BOUNCE, rr_depth=RR, threads=cores)
            state["rays"] = int(st[0] + st[1])
            return time.perf_counter() - t0, {}
        times, _ = reference_steps(150.0, max(1, args.steps), run)
        ms = 1e3 * float(np.min(times))
        mrays = state["rays"] / (ms * 1e3)
        msamples = REF_W * REF_H * REF_SPP / (ms * 1e3)
    sample = (f"{SCENE} {REF_W}x{REF_H} x {REF_SPP} spp x {BOUNCE} bounces per step (the bench frame at 1 of its 16 samples), "
              f"{cores} OpenMP threads, best of {len(times)}")
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": round(mrays, 4), "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": len(times), "warmup": min(args.warmup, 1), "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "msamples_per_s": round(msamples, 4),
        "config": {"workload": WORKLOAD, "sample": sample, "width": REF_W, "height": REF_H, "spp": REF_SPP, "bounces": BOUNCE},
        "cpu_baseline": {"value": round(mrays, 4), "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(mrays, 4), "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def cpu_baseline_leg():
    """Bounded sample of the same workload on the host cores (rank 0, N = 1): the reference itself
    when oracle/_ref travelled with the repo, else the oracle port.  Also returns the traversal
    counters on the reference layout that define the algorithmic bytes per ray."""
    from tests import oracle_lib as ol
    from aten_b200 import scene_io as sio
    cores = os.cpu_count() or 1
    sw, sh = 480, 270
    scene = sio.load_scene(os.path.join(ROOT, "tests", "golden", f"{SCENE}.atnscene.gz")).with_size(sw, sh)
    t0 = time.perf_counter()
    _, _, st = ol.render(scene, frames=1, max_depth=BOUNCE, rr_depth=RR, threads=cores)
    port_s = time.perf_counter() - t0
    rays = float(st[0] + st[1])
    n_node, n_box, n_tri, n_inst = (float(st[i]) / rays for i in (2, 3, 4, 5))
    # B_ray = 24 (ray in) + 32 (hit out) + 48 N_node + 80 N_tri + 64 N_inst + 64 N_xform  
    b_ray = 24 + 32 + 48 * n_node + 80 * n_tri + 64 * n_inst + 64 * n_inst
    f_ray = 34 * n_box + 58 * n_tri + 30 * n_inst
    out = {"b_ray": b_ray, "f_ray": f_ray, "n_node": n_node, "n_box": n_box, "n_tri": n_tri, "n_inst": n_inst}
    if ol.have_ref():
        info = ol.ref_run("bench", SCENE, REF_W, REF_H, REF_SPP, BOUNCE, RR, cores, 1, timeout=1800)
        out["cpu"] = {"value": round(info["mrays_per_s"], 4), "unit": "Mrays/s", "cores": cores, "kind": "reference",
                      "sample": f"{SCENE} {REF_W}x{REF_H} x {REF_SPP} spp x {BOUNCE} bounces (the bench frame at 1 of its 16 samples), "
                                f"aten::PathTracing::render, {cores} OpenMP threads, one run after warm-up",
                      "msamples_per_s": round(info["msamples_per_s"], 4)}
    else:
        out["cpu"] = {"value": round(rays / port_s / 1e6, 4), "unit": "Mrays/s", "cores": cores, "kind": "port",
                      "sample": f"{SCENE} {sw}x{sh} x 1 spp x {BOUNCE} bounces, oracle/pt_oracle.cpp, {cores} OpenMP threads"}
    return out


def extra_workloads():
    """The heavier stand-in SURVEY 8(d) asks for next to configs[1], and configs[4] -- the one config whose working set is
    far beyond the L2, where an HBM fraction means something.  Needs the reference's loaders / SBVH builder (oracle/_ref)."""
    from tests import oracle_lib as ol
    if not ol.have_ref():
        return {"unavailable": "oracle/_ref did not travel (scenes are built by the reference's loaders and SBVH builder)"}
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import bench_configs as bc
    out = {}
    t0 = time.perf_counter()
    out["instanced_dragon_1080p_16spp"] = bc.render_bench(
        "instanced_dragon", 1920, 1080, 16, steps=3, warmup=2,
        label="heavier configs[1] / configs[3] stand-in: instanced dragon (100 k triangles) + teapots + suzannes, Disney / refraction / GGX")
    c5 = bc.ray_stress_flat("dragon_grid10", 100_000_000, render=False,
                            label="configs[4]: 100 M incoherent secondary rays into 100 distinct dragon meshes (10 M triangles, 1.3 GB of nodes + triangles)")
    out["c5_10M_triangles_100M_rays"] = c5
    out["seconds"] = round(time.perf_counter() - t0, 1)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--partition", default="samples", choices=["samples", "tiles"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the other workloads (heavier stand-in, configs[4])")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from aten_b200 import PathTracing, multigpu, scene_io as sio
    from aten_b200.api import GATHER_SUM, probe_read_bandwidth

    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device; aten_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene = sio.load_scene(os.path.join(ROOT, "tests", "golden", f"{SCENE}.atnscene.gz")).with_size(W, H)
    r = PathTracing(local_rank)
    seeds = sio.mt19937_seeds(W * H)
    r.UpdateSceneData(scene, seeds=seeds)
    r.set_gather_mode(GATHER_SUM)         # sum + count in .w: the form the reduce combines

    img_ptr, npix = r.device_image()

    class _Img:  # zero-copy torch view of the renderer's float4 image
        __cuda_array_interface__ = {"shape": (H, W, 4), "typestr": "<f4", "data": (img_ptr, False), "version": 3}
    image_t = torch.as_tensor(_Img(), device=f"cuda:{local_rank}")
    pinned_img = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
    pinned_seeds = torch.from_numpy(seeds.view(np.int32)).pin_memory()
    # everything of a step -- clear, render, reduce -- is ordered on the renderer's own stream
    rstream = torch.cuda.ExternalStream(r.stream(), device=f"cuda:{local_rank}")

    # this rank's share of the frame
    if world > 1 and args.scaling == "weak":
        s_begin, s_end = rank * SPP, rank * SPP + SPP              # its own 16 samples of a 16 N spp frame
        tiles = None
    elif args.partition == "tiles":
        s_begin, s_end = 0, SPP
        tiles = multigpu.partition_tiles(W, H, world, rank, 64) if world > 1 else None
    else:
        s_begin, s_end = multigpu.partition_samples(SPP, world, rank)
        tiles = None
    total_spp = SPP * world if (world > 1 and args.scaling == "weak") else SPP

    def render_share():
        r.reset()
        if tiles is not None:
            r.render_tiles(W, H, 64, world, rank, s_begin, s_end, BOUNCE, RR)
        elif s_end > s_begin:
            r.render_region(W, H, 0, 0, W, H, s_begin, s_end, BOUNCE, RR)

    def step_device():
        render_share()
        if world > 1:
            with torch.cuda.stream(rstream):
                dist.reduce(image_t, dst=0, op=dist.ReduceOp.SUM)      # ordered behind the render, ahead of the next clear

    def step_e2e():
        r.set_seeds(pinned_seeds.numpy().view(np.uint32))        # H2D from pinned host memory
        step_device()
        if rank == 0:
            r.readback_into(pinned_img.data_ptr())               # D2H of the step's result (waits for the stream)
        else:
            r.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    # ---- timed: device-resident, CUDA events on the renderer's stream, no host sync inside ----
    clocks = ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    clocks.start()
    t0 = time.perf_counter()
    e0.record(rstream)
    for _ in range(args.steps):
        step_device()
    e1.record(rstream)
    barrier()
    wall_s = time.perf_counter() - t0
    clk = clocks.stop()
    dev_ms = e0.elapsed_time(e1)
    st = r.stats()                        # the work of one step (every step renders the same samples)
    rays = (st["closest_rays"] + st["shadow_rays"]) * args.steps
    samples = st["camera_samples"] * args.steps
    launches = st["kernel_launches"] * args.steps
    nodes = tris = 0                      # counted in the profiling pass below (the counting kernels are not the timed ones)
    exact = st["exact_rays"] * args.steps

    # ---- timed: end to end through the public API with host buffers ----
    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    e_rays = rays

    # ---- traversal kernels alone (CUDA events around every traversal launch) ----
    tr_ms = tr_rays = 0.0
    trace_launches = 0
    prof_ms = 0.0
    if s_end > s_begin and tiles is None:
        r.set_profiling(True)
        render_share(); r.stats()
        for _ in range(max(1, min(args.steps, 3))):
            render_share()
            ps = r.stats()
            tr_ms += ps["trace_ms"]; prof_ms += ps["device_ms"]
            tr_rays += ps["closest_rays"] + ps["shadow_rays"]
            trace_launches += 2 * BOUNCE * (-(-(s_end - s_begin) // 16))     # up to 16 samples of a 1080p frame per wavefront
        r.set_profiling(2)                # node fetches / triangle tests per step: the counting form of the kernels, untimed
        render_share()
        cs = r.stats()
        nodes = cs["nodes_visited"] * args.steps; tris = cs["tris_tested"] * args.steps
        r.set_profiling(False)
    barrier()

    # ---- aggregate over ranks (max time, sum of work) ----
    if world > 1:
        t = torch.tensor([wall_s, e2e_s, dev_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall_s, e2e_s, dev_ms = (float(x) for x in t.tolist())
        w = torch.tensor([rays, samples, launches, e_rays, nodes, tris, exact], dtype=torch.float64, device="cuda")
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
        rays, samples, launches, e_rays, nodes, tris, exact = (float(x) for x in w.tolist())

    if rank == 0:
        hbm, which = peaks()
        base = {"b_ray": None}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            base = cpu_baseline_leg()
            cpu = base["cpu"]
        # throughput over the timed region: CUDA-event time on the renderer's stream, max over ranks (with N > 1 it
        # includes the NCCL reduce, which runs on that stream)
        t_s = dev_ms / 1e3
        value = rays / t_s / 1e6
        if world == 1:
            part = "single GPU"
        elif args.scaling == "weak":
            part = f"weak scaling: rank r renders its own {SPP} samples (a {total_spp} spp frame); NCCL reduce(sum) of float4 images to rank 0 on the render stream"
        elif args.partition == "tiles":
            part = f"strong scaling of the fixed {SPP} spp frame by interleaved 64x64 tiles; NCCL reduce(sum) of float4 images to rank 0 on the render stream"
        else:
            part = (f"strong scaling of the fixed {SPP} spp frame by sample index: rank r renders {SPP}/{world} of every pixel's samples; "
                    "NCCL reduce(sum) of float4 images to rank 0 on the render stream")
        line = {
            "metric": "Mrays/s", "value": round(value, 2), "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(1e3 * t_s / args.steps, 3), "higher_is_better": True,
            "scaling": "weak" if (world > 1 and args.scaling == "weak") else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "msamples_per_s": round(samples / t_s / 1e6, 2),
            "rays_per_sample": round(rays / max(samples, 1), 3),
            "own_layout": {"quad_nodes_per_ray": round(nodes / max(rays, 1), 2), "tri_tests_per_ray": round(tris / max(rays, 1), 2),
                           "bytes_per_ray": round((128 * nodes + 64 * tris) / max(rays, 1) + 56, 1)},
            "exact_walk_share_of_rays": round(exact / max(rays, 1), 5),
            "wall_ms_per_step": round(1e3 * wall_s / args.steps, 3),
            "config": {"workload": WORKLOAD, "scene": SCENE, "width": W, "height": H, "spp": total_spp, "spp_per_gpu": s_end - s_begin,
                       "bounces": BOUNCE, "rr_bounce": RR, "partition": part,
                       "l2": "path state (33M paths per wavefront x ~250 B SoA = 8 GB touched per bounce) exceeds the 126 MB L2; "
                             "the scene (1.2 MB nodes + 1.2 MB triangles) is L2-resident by design, as in production use"},
            "clocks": clk,
            "e2e": {"value": round(e_rays / e2e_s / 1e6, 2), "unit": "Mrays/s",
                    "h2d_bytes_per_step": int(seeds.nbytes), "d2h_bytes_per_step": int(W * H * 16),
                    "ms_per_step": round(1e3 * e2e_s / args.steps, 3)},
            "gpu_launches": int(launches),
        }
        if base.get("b_ray") and tr_ms > 0:
            ncu = json.load(open(NCU_SUMMARY)) if os.path.exists(NCU_SUMMARY) else {}
            sm_mhz = clk.get("sm_mhz") or 1965.0
            lane_peak = 148 * 4 * 32 * sm_mhz * 1e6 / 1e12         # T lane-instructions / s the SMs can issue
            issue = ncu.get("issue_slot_utilisation"); tpi = ncu.get("threads_per_instruction")
            frac = round(issue * tpi / 32.0, 4) if (issue and tpi) else None
            alg_gbs = tr_rays * base["b_ray"] / (tr_ms / 1e3) / 1e9
            own_bytes = line["own_layout"]["bytes_per_ray"]
            line["roofline"] = {
                # what binds the traversal kernels: instruction issue with partly empty warps
                "bound": "issue", "unit": "T lane-instructions/s",
                "achieved": round(frac * lane_peak, 2) if frac else None, "peak": round(lane_peak, 2), "frac": frac,
                "issue_slot_utilisation_ncu": issue, "threads_per_instruction_ncu": tpi,
                "source": ncu.get("source", "no ncu summary committed for this build"),
                "kernel": "k_trace<ClosestPolicy> + k_trace<ShadowPolicy>",
                "trace_share_of_step": round(tr_ms / max(1e-9, prof_ms), 3),
                "trace_mrays_per_s": round(tr_rays / (tr_ms / 1e3) / 1e6, 1),
                # dram__bytes_read + dram__bytes_write per traversal launch (ncu --set full): the ray / hit streams only
                "traffic": ncu.get("dram_bytes_per_trace_launch"),
                "algorithmic_bytes_per_launch": round(tr_rays * base["b_ray"] / max(1, trace_launches), 0),
                "trace_launches_per_step": trace_launches // max(1, min(args.steps, 3)),
                # SURVEY 8(d)'s prescribed figure: algorithmic bytes on the REFERENCE layout and visit order (oracle counters)
                # per second of traversal-kernel time against the HBM copy peak.  It measures the algorithmic saving of the
                # ordered 4-slot walk over the reference's unordered binary walk on an L2-resident set -- not DRAM efficiency
                "reference_layout": {"bytes_per_ray": round(base["b_ray"], 1), "flops_per_ray": round(base["f_ray"], 1),
                                     "nodes_per_ray": round(base["n_node"], 2), "tris_per_ray": round(base["n_tri"], 2),
                                     "algorithmic_gbs": round(alg_gbs, 1), "hbm_copy_peak_gbs": hbm, "peak_source": which,
                                     "ratio_to_hbm_copy_peak": round(alg_gbs / hbm, 3),
                                     "algorithmic_tflops": round(tr_rays * base["f_ray"] / (tr_ms / 1e3) / 1e12, 2),
                                     "fp32_peak_tflops": round(148 * 128 * 2 * sm_mhz * 1e6 / 1e12, 1)},
                # the L2 level: bytes the traversal launches move through the L2 (lts__t_bytes, ncu) per second of their own
                # time, against the read bandwidth one persistent launch gets from an L2-resident buffer (measured below)
                "l2": None,
            }
            if world == 1:
                try:
                    l2_gbs, l2_ms = probe_read_bandwidth(local_rank, 32 << 20, 4000)
                    hbm_rd, _ = probe_read_bandwidth(local_rank, 4 << 30, 8)
                    lts = ncu.get("lts_bytes_per_trace_launch")
                    trace_ms_per_launch = tr_ms / max(1, trace_launches)
                    line["roofline"]["l2"] = {
                        "probe_read_gbs_32MB_resident": round(l2_gbs, 1), "probe_ms": round(l2_ms, 2),
                        "probe_read_gbs_4GB_from_hbm": round(hbm_rd, 1),
                        "lts_bytes_per_trace_launch_ncu": lts,
                        "achieved_gbs": round(lts / (trace_ms_per_launch * 1e-3) / 1e9, 1) if lts else None,
                        "frac": round(lts / (trace_ms_per_launch * 1e-3) / 1e9 / l2_gbs, 3) if lts else None,
                        "requested_at_l1_gbs": round(tr_rays * own_bytes / (tr_ms / 1e3) / 1e9, 1),
                    }
                except Exception as ex:        # the probe is an aid, never a reason to lose the line
                    line["roofline"]["l2"] = {"error": str(ex)}
        if cpu:
            line["cpu_baseline"] = cpu
        if world == 1 and not args.no_extra:
            try:
                r.close(); r = None
                line["other_workloads"] = extra_workloads()
            except Exception as ex:
                line["other_workloads"] = {"error": str(ex)[:300]}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if r is not None:
        r.close()


if __name__ == "__main__":
    main()
