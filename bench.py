#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on its config: Mrays/s (closest-hit + shadow rays actually
traced) and Msamples/s for "Sponza 1920x1080, 16 spp, 5 bounces" on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W            # our arm
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle/_ref)

A step = one pass of the hot path over one batch: one atn_render of the whole 1920x1080 frame at
16 spp with 5 bounces (genPath -> [closest-hit, shade, shadow-ray] x 5 -> gather, 16 times).
Scene: the `sponza_atrium` blob (asset/sponza/sponza_lod.obj + sponza_lod.sbvh, Lambert/GGX, area
light + synthetic-sky IBL; the full-resolution Sponza OBJ is stripped from the reference snapshot,
SURVEY.md F10).  Synthetic data: seeds = std::mt19937(0), camera fixed.

value  : device-resident throughput; scene, seeds and path state already in HBM, CUDA-event time.
e2e    : same metric through the public API with HOST buffers: per step the per-pixel seeds are
         copied host->device from pinned memory and the float4 image is read back to pinned memory.
N > 1  : one process per GPU (torchrun), scene replicated, rank r renders samples [16r, 16r+16) of the
         frame sequence (sample-index partition, weak scaling), float4 sum images combined with one
         NCCL reduce to rank 0 inside the timed region.
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
L2_COPY_PEAK_GBS = 8140.2   # measured on this pool's B200: copy between two L2-resident 32 MB buffers (profiles/r01_l2_bandwidth.json)
WORKLOAD = "sponza_lod 1920x1080 16spp 5-bounce NEE (configs[1] stand-in: full Sponza OBJ missing from snapshot)"


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


def reference_arm(args, rank: int, world: int):
    """The reference's own CPU path (aten::PathTracing over AcceleratedScene<sbvh>, compiled in place
    into oracle/_ref) on the box's host cores.  Each step is a BOUNDED sample of the workload: the
    same scene / camera / seeds / bounces at 960x540, 1 spp."""
    if rank != 0:
        return
    from tests import oracle_lib as ol
    cores = os.cpu_count() or 1
    sw, sh, frames = 960, 540, 1
    if not ol.have_ref():
        # oracle port as the stand-in (kind "port")
        from aten_b200 import scene_io as sio
        scene = sio.load_scene(os.path.join(ROOT, "tests", "golden", f"{SCENE}.atnscene.gz")).with_size(sw, sh)
        times, rays = [], 0
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            _, _, st = ol.render(scene, frames=frames, max_depth=BOUNCE, rr_depth=RR, threads=cores)
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
            rays = int(st[0] + st[1])
        kind = "port"
        ms = 1e3 * float(np.mean(times))
        mrays = rays / (ms * 1e3)
        msamples = sw * sh * frames / (ms * 1e3)
    else:
        info = ol.ref_run("bench", SCENE, sw, sh, frames, BOUNCE, RR, cores, max(1, args.steps), timeout=3000)
        kind = "reference"
        ms = info["best_ms"]
        mrays = info["mrays_per_s"]
        msamples = info["msamples_per_s"]
    sample = f"{SCENE} {sw}x{sh} x {frames} spp x {BOUNCE} bounces per step, {cores} OpenMP threads, best of {max(1, args.steps)}"
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": round(mrays, 4), "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "msamples_per_s": round(msamples, 4),
        "config": {"workload": WORKLOAD, "sample": sample},
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
        info = ol.ref_run("bench", SCENE, 960, 540, 1, BOUNCE, RR, cores, 1, timeout=1800)
        out["cpu"] = {"value": round(info["mrays_per_s"], 4), "unit": "Mrays/s", "cores": cores, "kind": "reference",
                      "sample": f"{SCENE} 960x540 x 1 spp x {BOUNCE} bounces, aten::PathTracing::render, {cores} OpenMP threads, best of 1 after warm-up",
                      "msamples_per_s": round(info["msamples_per_s"], 4)}
    else:
        out["cpu"] = {"value": round(rays / port_s / 1e6, 4), "unit": "Mrays/s", "cores": cores, "kind": "port",
                      "sample": f"{SCENE} {sw}x{sh} x 1 spp x {BOUNCE} bounces, oracle/pt_oracle.cpp, {cores} OpenMP threads"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from aten_b200 import PathTracing, scene_io as sio

    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device; aten_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene = sio.load_scene(os.path.join(ROOT, "tests", "golden", f"{SCENE}.atnscene.gz")).with_size(W, H)
    r = PathTracing(local_rank)
    seeds = sio.mt19937_seeds(W * H)
    r.UpdateSceneData(scene, seeds=seeds)
    r.set_gather_mode(2)         # sum + count in .w: the form the reduce combines

    img_ptr, npix = r.device_image()

    class _Img:  # zero-copy torch view of the renderer's float4 image
        __cuda_array_interface__ = {"shape": (H, W, 4), "typestr": "<f4", "data": (img_ptr, False), "version": 3}
    image_t = torch.as_tensor(_Img(), device=f"cuda:{local_rank}")
    pinned_img = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
    pinned_seeds = torch.from_numpy(seeds.view(np.int32)).pin_memory()
    sample_base = rank * SPP                       # sample-index partition

    def step_device():
        r.reset()
        r.render_region(W, H, 0, 0, W, H, sample_base, sample_base + SPP, BOUNCE, RR)
        st = r.stats()
        if world > 1:
            dist.reduce(image_t, dst=0, op=dist.ReduceOp.SUM)
            # the reduce runs on torch's stream, the next step clears the image on the renderer's: finish it first
            torch.cuda.current_stream().synchronize()
        return st

    def step_e2e():
        r.set_seeds(pinned_seeds.numpy().view(np.uint32))        # H2D from pinned host memory
        r.reset()
        r.render_region(W, H, 0, 0, W, H, sample_base, sample_base + SPP, BOUNCE, RR)
        st = r.stats()
        if world > 1:
            dist.reduce(image_t, dst=0, op=dist.ReduceOp.SUM)
            torch.cuda.synchronize()
        if rank == 0:
            r.readback_into(pinned_img.data_ptr())               # D2H of the step's result
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    # ---- timed: device-resident ----
    clocks = ClockSampler(local_rank)
    barrier()
    clocks.start()
    t0 = time.perf_counter()
    rays = samples = launches = nodes = tris = 0
    dev_ms = 0.0
    for _ in range(args.steps):
        st = step_device()
        nodes += st["nodes_visited"]; tris += st["tris_tested"]
        rays += st["closest_rays"] + st["shadow_rays"]
        samples += st["camera_samples"]
        launches += st["kernel_launches"]
        dev_ms += st["device_ms"]
    barrier()
    wall_s = time.perf_counter() - t0
    clk = clocks.stop()

    # ---- timed: end to end through the public API with host buffers ----
    step_e2e()
    barrier()
    t0 = time.perf_counter()
    e_rays = 0
    for _ in range(args.steps):
        st = step_e2e()
        e_rays += st["closest_rays"] + st["shadow_rays"]
    barrier()
    e2e_s = time.perf_counter() - t0

    # ---- traversal kernels alone (CUDA events around every traversal launch) ----
    r.set_profiling(True)
    step_device()
    tr_ms = tr_rays = 0.0
    trace_launches = 0
    for _ in range(max(1, min(args.steps, 3))):
        st = step_device()
        tr_ms += st["trace_ms"]
        tr_rays += st["closest_rays"] + st["shadow_rays"]
        trace_launches += 2 * BOUNCE * SPP
    r.set_profiling(False)
    barrier()

    # ---- aggregate over ranks (max time, sum of work) ----
    if world > 1:
        t = torch.tensor([wall_s, e2e_s, dev_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall_s, e2e_s, dev_ms = (float(x) for x in t.tolist())
        w = torch.tensor([rays, samples, launches, e_rays], dtype=torch.float64, device="cuda")
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
        rays, samples, launches, e_rays = (float(x) for x in w.tolist())

    if rank == 0:
        hbm, which = peaks()
        base = {"b_ray": None}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            base = cpu_baseline_leg()
            cpu = base["cpu"]
        # throughput over the timed region: with N = 1 the CUDA-event time of the renders; with N > 1
        # the barrier-to-barrier wall clock (includes the NCCL reduce), max over ranks
        t_s = (dev_ms / 1e3) if world == 1 else wall_s
        value = rays / t_s / 1e6
        line = {
            "metric": "Mrays/s", "value": round(value, 2), "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(1e3 * t_s / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "msamples_per_s": round(samples / t_s / 1e6, 2),
            "rays_per_sample": round(rays / max(samples, 1), 3),
            "own_layout": {"quad_nodes_per_ray": round(nodes * world / max(rays, 1), 2), "tri_tests_per_ray": round(tris * world / max(rays, 1), 2),
                           "bytes_per_ray": round((128 * nodes + 64 * tris) * world / max(rays, 1) + 56, 1)},
            "wall_ms_per_step": round(1e3 * wall_s / args.steps, 3),
            "config": {"workload": WORKLOAD, "scene": SCENE, "width": W, "height": H, "spp_per_gpu": SPP,
                       "bounces": BOUNCE, "rr_bounce": RR,
                       "partition": "sample index: rank r renders samples [16r,16r+16); NCCL reduce(sum) of float4 images to rank 0" if world > 1 else "single GPU",
                       "l2": "path state (8.3M paths per wavefront x ~200 B SoA = 1.7 GB touched per bounce) exceeds the 126 MB L2; "
                             "the scene (2.4 MB nodes + 0.6 MB triangles) is L2-resident by design, as in production use"},
            "clocks": clk,
            "e2e": {"value": round(e_rays / e2e_s / 1e6, 2), "unit": "Mrays/s",
                    "h2d_bytes_per_step": int(seeds.nbytes), "d2h_bytes_per_step": int(W * H * 16),
                    "ms_per_step": round(1e3 * e2e_s / args.steps, 3)},
            "gpu_launches": int(launches),
        }
        if base.get("b_ray"):
            achieved = tr_rays * base["b_ray"] / (tr_ms / 1e3) / 1e9
            line["roofline"] = {
                "bound": "hbm", "achieved": round(achieved, 1), "peak": hbm, "unit": "GB/s", "frac": round(achieved / hbm, 4),
                # dram__bytes_read+write per traversal launch from the ncu --set full capture summarised in
                # profiles/r01_ncu_full_v9_batched.txt (0.47-0.57 GB per launch of a four-sample wavefront, mean of the nine
                # full launches captured): the ray / hit streams only -- nodes and triangles are served from L2
                "traffic": 5.1e8, "algorithmic_bytes_per_launch": round(tr_rays * base["b_ray"] / max(1, trace_launches), 0),
                "peak_source": which, "kernel": "k_trace<ClosestPolicy> + k_trace<ShadowPolicy>",
                "issue_slot_utilisation_ncu": 0.72, "threads_per_instruction_ncu": 15.9, "l1_throughput_ncu": 0.68,
                "bytes_per_ray": round(base["b_ray"], 1), "flops_per_ray": round(base["f_ray"], 1),
                # the second roof of SURVEY 8d: algorithmic flops against the FP32 peak of the SMs (148 x 128 lanes x 2 x the
                # SM clock sampled during the run); nominal, not measured
                "achieved_tflops": round(tr_rays * base["f_ray"] / (tr_ms / 1e3) / 1e12, 2),
                "fp32_peak_tflops": round(148 * 128 * 2 * (clk.get("sm_mhz") or 1965.0) * 1e6 / 1e12, 1),
                "fp32_frac": round(tr_rays * base["f_ray"] / (tr_ms / 1e3) / (148 * 128 * 2 * (clk.get("sm_mhz") or 1965.0) * 1e6), 4),
                "nodes_per_ray": round(base["n_node"], 2), "tris_per_ray": round(base["n_tri"], 2),
                "trace_share_of_step": round(tr_ms / max(1e-9, (dev_ms / args.steps) * max(1, min(args.steps, 3))), 3),
                "trace_mrays_per_s": round(tr_rays / (tr_ms / 1e3) / 1e6, 1),
                # the same kernels on their OWN layout: bytes they request (128 B per QuadNode step, 64 B per triangle test,
                # 56 B ray in / hit out; counted by the kernels) per second, against the L2-resident copy bandwidth of this
                # GPU (32 MB buffers, scripts/l2_bandwidth.py -> profiles/r01_l2_bandwidth.json) -- the roof that applies to
                # a scene served from L2
                "own_layout_fetched_gbs": round(tr_rays * line["own_layout"]["bytes_per_ray"] / (tr_ms / 1e3) / 1e9, 1),
                "l2_copy_peak_gbs": L2_COPY_PEAK_GBS,
                "own_layout_frac_of_l2_copy_peak": round(tr_rays * line["own_layout"]["bytes_per_ray"] / (tr_ms / 1e3) / 1e9 / L2_COPY_PEAK_GBS, 3),
                "note": "algorithmic bytes on the REFERENCE layout and visit order (oracle counters, SURVEY 8d); the scene is L2-resident and "
                        "the ordered traversal over 4-slot nodes fetches ~2.5 KB/ray from L1/L2 (own_layout), so frac > 1 against the HBM copy peak is expected; "
                        "the kernel is instruction-issue bound (issue_slot_utilisation_ncu at threads_per_instruction_ncu of 32 lanes); "
                        "part of every node is fetched through the texture pipe, which took l1tex throughput from 0.79 to l1_throughput_ncu",
            }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    r.close()


if __name__ == "__main__":
    main()
