#!/usr/bin/env python3
"""bench.py — Mrays/s of the ray-traversal hot path on B200 (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scene world16] ...

A step is one frame = BVH rebuild + trace, the scope of the reference's bench()
(/root/reference/src/main.cc:210-216 -> rtracer::gpu::update_scene, src/raytracer.cu:102-120).
Default workload: world16.json at 3840x2160, depth 9 (BASELINE.json configs[3], the headline).

  value : frame with everything resident in HBM (canvas stays on the device)
  e2e   : the same frame through the C ABI with HOST buffers — instances and camera uploaded,
          canvas copied back into pinned host memory — inside the timed region
  N > 1 : screen tiles dealt round-robin to the ranks, scene + BVH replicated; the frame's kernels
          store their pixels straight into the gathering rank's canvas over NVLink peer memory and
          signal through a mailbox (RT_RENDER_PEER_STORE, --gather fused-root; fused-all = every
          rank receives the frame). p2p / nccl: the round-1 exchanges (slab + push kernel, NCCL
          all-gather + de-interleave), kept as checked alternatives ("strong" scaling)
  ref_cuda / stress : N = 1 — the reference's own CUDA build (oracle/_ref/ref_gpu, bench() scope)
          timed on the same GPU in the same run, and the same measurement on world8_stress

`--impl reference` times the reference's own serial CPU path (oracle/_ref/ref_cpu, built from the
unmodified sources; falls back to the oracle port) on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s (all rays: primary + shadow + secondary closest-hit queries)"
UNIT = "Mrays/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="world16")
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--depth", type=int, default=9)
    ap.add_argument("--grid-size", type=int, default=0)
    ap.add_argument("--tile-w", type=int, default=32)
    ap.add_argument("--tile-h", type=int, default=16)
    ap.add_argument("--gather", default="fused-root", choices=["fused-root", "fused-all", "p2p", "p2p-nccl", "nccl"],
                    help="N > 1, device-resident frame: fused-root = the kernels that produce the pixels store them into rank 0's "
                         "canvas over NVLink peer memory and signal its mailbox (the framebuffer gathered, no extra kernel, no "
                         "collective); fused-all = the same into every rank's canvas; p2p = slab + rt_push_tiles_signal into every "
                         "rank's canvas; p2p-nccl = rt_push_tiles + 1-element all-reduce; nccl = all-gather of the slabs + "
                         "de-interleave kernel")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip the reference-CUDA-build leg (N = 1)")
    ap.add_argument("--no-fast-accept", action="store_true", help="skip the tolerance-mode block (N = 1)")
    ap.add_argument("--no-stress", action="store_true", help="skip the world8_stress block (N = 1)")
    ap.add_argument("--split-shade", action="store_true", help="RT_RENDER_SPLIT_SHADE: round 1's split pipeline (per-lane k_primary)")
    ap.add_argument("--graph", action="store_true", help="RT_RENDER_GRAPH: device-resident frames captured once and replayed as one CUDA graph")
    ap.add_argument("--lanes", type=int, default=0,
                    help="rt_scene_set_lanes: concurrent lanes of the device-resident frame (0 = library default, i.e. RT_FRAME_LANES or one stream; measured neutral, profiles/r2_lanes_ab.txt)")
    ap.add_argument("--tile-kernel", action="store_true", help="RT_RENDER_TILE_KERNEL: wave 0 entirely in the tile kernels")
    ap.add_argument("--band-h", type=int, default=4,
                    help="N > 1, e2e: rows per full-width band dealt round-robin to the ranks (each rank copies its bands into the one shared host canvas)")
    ap.add_argument("--host-chunks", type=int, default=0, help="e2e: chunks per frame for copy/render overlap (0 = the library's default)")
    ap.add_argument("--host-direct", action="store_true",
                    help="e2e: let the kernels write the pinned host canvas over PCIe instead of rendering into a device canvas and copying it (measured slower)")
    ap.add_argument("--sort-rays", action="store_true", help="RT_RENDER_SORT_RAYS: sort the live paths between bounce waves")
    ap.add_argument("--keep-bvh", action="store_true", help="do not rebuild the BVH every frame")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work budget of the cpu_baseline leg")
    return ap.parse_args()


def workload_name(a):
    return "%s.json %dx%d depth %d%s" % (a.scene, a.width, a.height, a.depth,
                                          (" grid %d" % a.grid_size) if a.grid_size else "")


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock and throttle reasons DURING the timed region: the nvidia-smi query loop of
    B200_PROFILING.md in a separate process (an in-process NVML thread was measured to add 2.5 ms to
    every multi-rank step), spawned before warm-up so that it is already streaming when the timed
    region starts. Samples are tagged with their arrival time and only those inside the region kept."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.t0, self.t1, self.proc = [], None, None, None
        period = os.environ.get("BENCH_CLOCK_PERIOD_MS", "10")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", period],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def wait_ready(self, timeout=15.0):
        """Blocks until nvidia-smi has printed its first row (its start-up takes seconds on an 8-GPU box; a
        multi-rank timed region of a few tens of milliseconds used to be over before the first sample)."""
        t_end = time.time() + timeout
        while self.proc and not self.rows and self.proc.poll() is None and time.time() < t_end:
            time.sleep(0.02)

    def start(self):
        self.t0 = time.time()

    def stop(self):
        self.t1 = time.time()
        time.sleep(0.05)
        if self.proc:
            self.proc.terminate()
        n_all = len(self.rows)
        # a region shorter than a few sampling periods (8 GPUs: 20 steps last ~15 ms) may hold no sample:
        # widen the window into the warm-up before it and the e2e leg after it (same workload, same load)
        for pad in (0.0, 0.1, 0.5):
            sm, mx, reasons = [], [], set()
            for t, r in self.rows:
                if self.t0 is None or not (self.t0 - pad <= t <= self.t1 + 0.03 + pad):
                    continue
                try:
                    sm.append(float(r[0])); mx.append(float(r[1]))
                    for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                        if v.lower().startswith("active"):
                            reasons.add(name)
                except Exception:
                    pass
            if len(sm) >= 2:
                break
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "samples_outside_region": n_all}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "region_ms": 1e3 * (self.t1 - self.t0), "window_pad_ms": 1e3 * pad}


# ------------------------------------------------------------------ reference CPU arm
def cpu_reference_run(a, budget_s, frames=1):
    """Times the reference's serial path on a bounded sample of the workload: the same scene and
    depth at a reduced resolution (rows/cols scaled by the same factor) so that one frame takes
    about budget_s/frames seconds. Returns (Mrays/s, kind, cores, sample, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import rtb200_loader
    pkg = rtb200_loader.load()
    import oracle_binding as ob
    ref_cpu = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")
    scene_json = os.path.join(ROOT, "scenes", a.scene + ".json")
    # ~1.0e5 rays/s/core on world16 for the reference's serial path; pick a
    # resolution whose ray count fits the budget, then measure — the estimate only sizes the sample
    est_rays_per_s = {"world16": 0.9e5, "config": 0.9e5, "world8_stress": 1.4e5, "world8": 3.0e5}.get(a.scene, 5e5)
    full_pixels = a.width * a.height
    frac = min(1.0, (est_rays_per_s * budget_s / frames) / (2.0 * full_pixels))
    scale = frac ** 0.5
    w = max(64, int(a.width * scale) // 8 * 8)
    h = max(48, int(a.height * scale) // 4 * 4)
    world = pkg.World(scene_json, width=w, height=h, depth=a.depth, grid_size=a.grid_size)
    # ray count of the sample = closest-hit queries of the serial path (counted by the oracle port)
    _, st = ob.OracleScene(world).render(semantics=pkg.RT_SEMANTICS_CPU, use_bvh=1)
    rays = st["rays_total"]
    if os.path.exists(ref_cpu):
        cfg = json.load(open(scene_json))
        cfg.update(width=w, height=h, depth=a.depth)
        if a.grid_size:
            cfg["grid_size"] = a.grid_size
        with tempfile.NamedTemporaryFile("w", suffix=".json", delete=False) as f:
            json.dump(cfg, f)
            tmp = f.name
        try:
            out = subprocess.run([ref_cpu, tmp, "--frames", str(frames)], capture_output=True, text=True, check=True).stdout
            info = json.loads(out.strip().splitlines()[-1])
            secs = info["ms_median"] * 1e-3
        finally:
            os.unlink(tmp)
        kind, cores = "reference", 1
    else:
        t0 = time.perf_counter()
        for _ in range(frames):
            ob.OracleScene(world).render(semantics=pkg.RT_SEMANTICS_CPU, use_bvh=1, threads=1)
        secs = (time.perf_counter() - t0) / frames
        kind, cores = "port", 1
    sample = "%s %dx%d depth %d, %d frame(s) of rtracer::cpu::update_scene (BVH build + trace), %d rays/frame, %.2f s/frame" % (
        a.scene, w, h, a.depth, frames, rays, secs)
    return rays / secs / 1e6, kind, cores, sample, secs


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    total = a.steps + a.warmup
    per_step = max(1.0, min(8.0, 150.0 / max(total, 1)))
    vals, kind, cores, sample = [], "port", 1, ""
    t_all = 0.0
    for i in range(total):
        v, kind, cores, sample, secs = cpu_reference_run(a, per_step, frames=1)
        if i >= a.warmup:
            vals.append(v)
            t_all += secs
    value = sum(vals) / len(vals)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * t_all / max(len(vals), 1), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic (procedural world, seed 42)",
            "config": {"workload": workload_name(a), "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
                             "host_cores": os.cpu_count()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0



# ------------------------------------------------------------------ the reference's own CUDA build, same GPU, same run
def ref_cuda_run(a, scene_name, frames=7):
    """oracle/_ref/ref_gpu (the unmodified reference compiled by oracle/ref_build, nvcc -arch=sm_100) on the bench
    workload: median ms of rtracer::gpu::update_scene(scene, 16, true) = BVH build + trace, the scope of its bench()
    (/root/reference/src/main.cc:65-68,210-216, -d 16). Its canvas lives in managed memory and is not read by the host
    inside the timed scope. Returns None when the binary is not there."""
    ref_gpu = os.path.join(ROOT, "oracle", "_ref", "ref_gpu")
    if not os.path.exists(ref_gpu):
        return None
    cfg = json.load(open(os.path.join(ROOT, "scenes", scene_name + ".json")))
    cfg.update(width=a.width, height=a.height, depth=a.depth)
    if a.grid_size:
        cfg["grid_size"] = a.grid_size
    with tempfile.NamedTemporaryFile("w", suffix=".json", delete=False) as f:
        json.dump(cfg, f)
        tmp = f.name
    try:
        out = subprocess.run([ref_gpu, tmp, "--gpu", "--frames", str(frames), "--kd", "16"], capture_output=True, text=True, timeout=600)
        if out.returncode != 0:
            return {"error": (out.stderr or out.stdout)[-200:]}
        return json.loads(out.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[-200:]}
    finally:
        os.unlink(tmp)


def ref_cuda_block(a, scene_name, rays_per_frame, ms_dev, ms_e2e):
    info = ref_cuda_run(a, scene_name)
    if info is None:
        return {"unavailable": "oracle/_ref/ref_gpu not built (needs /root/reference at build time)"}
    if "error" in info:
        return {"unavailable": info["error"]}
    ms = info["ms_median"]
    return {"ms": ms, "ms_min": info.get("ms_min"), "mrays_s": rays_per_frame / ms / 1e3,
            "scope": "rtracer::gpu::update_scene = BVH build + trace per frame, -d 16, canvas left in managed memory (the reference's bench() scope)",
            "binary": "oracle/_ref/ref_gpu (unmodified reference sources, nvcc -arch=sm_100), run on the same GPU after the timed region",
            "rays": "counted by this engine's frame of the same scene (both answer the same closest-hit queries)",
            "ratio_device_resident": ms / ms_dev, "ratio_e2e_pinned_host_canvas": ms / ms_e2e,
            "note": "ratio_device_resident compares like scopes (neither side reads the frame on the host); ratio_e2e charges this "
                    "engine the copy of the frame into pinned host memory, which the reference's scope does not contain"}


def single_gpu_block(pkg, torch, a, scene_name, steps, flush):
    """The N = 1 measurement of the main line (device-resident `value`, `e2e` with host buffers, reference CUDA build) for a
    second scene: world8_stress, the other scene north_star's >= 10x target names."""
    wfile = pkg.World(os.path.join(ROOT, "scenes", scene_name + ".json"), width=a.width, height=a.height, depth=a.depth)
    scene = pkg.Scene(wfile, device=int(os.environ.get("LOCAL_RANK", "0")))
    scene.set_stream(torch.cuda.current_stream().cuda_stream)
    W, H = wfile.width, wfile.height
    desc = wfile.desc
    canvas_dev = torch.zeros(H * W, dtype=torch.int32, device="cuda")
    host_canvas = torch.zeros(H * W, dtype=torch.int32).pin_memory()
    base = (0 if a.keep_bvh else pkg.RT_RENDER_REBUILD_BVH) | (pkg.RT_RENDER_SPLIT_SHADE if a.split_shade else 0) | \
        (pkg.RT_RENDER_TILE_KERNEL if a.tile_kernel else 0)

    def dev():
        scene.render(flags=base | pkg.RT_RENDER_CANVAS_DEVICE | pkg.RT_RENDER_ASYNC, out_ptr=canvas_dev.data_ptr())
        return None

    def e2e():
        scene.set_instances(desc.instances, desc.n_instances)
        scene.set_camera(desc.camera)
        _, st = scene.render(flags=base, out_ptr=host_canvas.data_ptr())
        return st

    def timed(fn, finish):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        total, outs = 0.0, []
        for _ in range(steps):
            flush.zero_()
            torch.cuda.synchronize()
            ev0.record()
            r = fn()
            ev1.record()
            ev1.synchronize()
            outs.append(r if r is not None else finish())
            total += ev0.elapsed_time(ev1)
        return total / steps, outs

    for _ in range(3):
        dev(); torch.cuda.synchronize(); scene.finish(); e2e()
    want = None
    ms_dev, outs = timed(dev, scene.finish)
    want = canvas_dev.cpu().numpy().view(np.uint32)
    ms_e2e, outs_e = timed(e2e, None)
    if not (host_canvas.numpy().view(np.uint32) == want).all():
        raise SystemExit("%s: end-to-end frame differs from the device-resident frame" % scene_name)
    rays = outs[0]["rays_total"]
    st = outs[0]
    block = {"workload": "%s.json %dx%d depth %d" % (scene_name, a.width, a.height, a.depth), "value": rays / ms_dev / 1e3, "unit": UNIT,
             "ms_per_step": ms_dev, "steps": steps, "rays_per_frame": rays, "rays_secondary": st["rays_secondary"], "waves": st["waves"],
             "e2e": {"value": rays / ms_e2e / 1e3, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": C.sizeof(pkg.RtInstance) * desc.n_instances + C.sizeof(pkg.RtCamera),
                     "d2h_bytes_per_step": 4 * W * H},
             "breakdown_ms": {k: sum(o[k] for o in outs) / len(outs) for k in ("ms_build", "ms_primary", "ms_shade", "ms_bounce")},
             "gpu_launches": sum(o["kernel_launches"] for o in outs)}
    if not a.no_ref_cuda:
        block["ref_cuda"] = ref_cuda_block(a, scene_name, rays, ms_dev, ms_e2e)
    scene.close()
    return block


# ------------------------------------------------------------------ our arm
def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def run_ours(a):
    import torch
    import torch.distributed as dist
    import rtb200_loader
    pkg = rtb200_loader.load()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    # everything of this rank (engine, torch kernels, NCCL, timing events) goes to ONE non-default stream: the legacy default
    # stream cannot be captured into a CUDA graph (--graph) and synchronises implicitly with every other stream
    torch.cuda.set_stream(torch.cuda.Stream())
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    wfile = pkg.World(os.path.join(ROOT, "scenes", a.scene + ".json"), width=a.width, height=a.height, depth=a.depth,
                      grid_size=a.grid_size)
    scene = pkg.Scene(wfile, device=local_rank)
    if a.lanes:
        scene.set_lanes(a.lanes)
    W, H = wfile.width, wfile.height
    base = (0 if a.keep_bvh else pkg.RT_RENDER_REBUILD_BVH) | (pkg.RT_RENDER_SORT_RAYS if a.sort_rays else 0) | \
        (pkg.RT_RENDER_SPLIT_SHADE if a.split_shade else 0) | (pkg.RT_RENDER_TILE_KERNEL if a.tile_kernel else 0)
    dev_flags = base | pkg.RT_RENDER_CANVAS_DEVICE | (pkg.RT_RENDER_TILED_OUTPUT if world > 1 else 0) | (pkg.RT_RENDER_GRAPH if a.graph else 0)
    params = scene.params(flags=dev_flags, rank=rank, world=world, tile_w=a.tile_w, tile_h=a.tile_h)
    canvas_dev = torch.zeros(H * W, dtype=torch.int32, device="cuda")
    host_canvas = torch.zeros(H * W, dtype=torch.int32).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    slab_t = gathered = None
    if world > 1:
        slab = scene.tiled_slab_pixels(params)
        slab_t = torch.zeros(slab, dtype=torch.int32, device="cuda")
        gathered = torch.zeros(world * slab, dtype=torch.int32, device="cuda")
    desc = wfile.desc
    h2d = C.sizeof(pkg.RtInstance) * desc.n_instances + C.sizeof(pkg.RtCamera)

    # the engine enqueues on torch's current stream: frame, NCCL all-gather and de-interleave are
    # stream-ordered, with no host synchronisation inside a step
    scene.set_stream(torch.cuda.current_stream().cuda_stream)
    params_async = scene.params(flags=dev_flags | pkg.RT_RENDER_ASYNC, rank=rank, world=world, tile_w=a.tile_w, tile_h=a.tile_h)

    # N > 1, default: no collective on the data path. Each rank owns a canvas the others map through
    # CUDA IPC and rt_push_tiles stores the rank's tiles into all of them; a one-element all-reduce on
    # the same stream is the barrier that tells every rank its canvas is complete.
    peer_ptrs, own_canvas_ptr, token = None, None, None
    epoch = [0]
    gather = a.gather if world > 1 else "none"
    fused = gather in ("fused-root", "fused-all")
    fused_sets = None  # [parity][rank] canvases of RT_RENDER_PEER_STORE
    if gather in ("p2p", "p2p-nccl") or fused:
        try:
            n_own = 2 if fused else 1
            own = [pkg.peer_canvas_alloc(W * H) for _ in range(n_own)]
            mine = torch.tensor([b for _, hnd in own for b in hnd], dtype=torch.uint8, device="cuda")
            allh = torch.empty(world * 64 * n_own, dtype=torch.uint8, device="cuda")
            dist.all_gather_into_tensor(allh, mine)
            allh = bytes(allh.cpu().numpy().tobytes())
            sets = [[own[k][0] if r == rank else pkg.peer_canvas_open(allh[64 * (r * n_own + k):64 * (r * n_own + k) + 64])
                     for r in range(world)] for k in range(n_own)]
            own_canvas_ptr = own[0][0]
            peer_ptrs = sets[0]
            if fused:
                fused_sets = sets
                scene.set_peers(sets[0] + sets[1], world, W * H, 0 if gather == "fused-root" else -1)
            ok = torch.ones(1, device="cuda")
        except RuntimeError as e:
            sys.stderr.write("peer canvases unavailable (%s); using the all-gather path\n" % e)
            ok = torch.zeros(1, device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok.item()) == 0.0:
            gather, peer_ptrs, fused, fused_sets = "nccl", None, False, None
        token = torch.zeros(1, device="cuda")
    fused_flags = (dev_flags & ~pkg.RT_RENDER_TILED_OUTPUT) | pkg.RT_RENDER_PEER_STORE

    def frame_device(flags=dev_flags):
        """One frame, canvas left in HBM (rank-local tiles gathered and de-interleaved when N > 1).
        Only enqueues; the caller synchronises and then collects the statistics with frame_stats()."""
        flags |= pkg.RT_RENDER_ASYNC
        if world == 1:
            scene.render(flags=flags, tile_w=a.tile_w, tile_h=a.tile_h, out_ptr=canvas_dev.data_ptr())
            return
        if fused:
            # the gather is part of the frame: k_tile stores into the peer canvases and signals; rank 0 (or every rank) waits
            scene.render(flags=(flags & ~pkg.RT_RENDER_TILED_OUTPUT) | pkg.RT_RENDER_PEER_STORE, rank=rank, world=world,
                         tile_w=a.tile_w, tile_h=a.tile_h, out_ptr=0)
            return
        scene.render(flags=flags, rank=rank, world=world, tile_w=a.tile_w, tile_h=a.tile_h, out_ptr=slab_t.data_ptr())
        if gather == "p2p":
            epoch[0] += 1
            scene.push_tiles_signal(params_async, slab_t.data_ptr(), peer_ptrs, W * H, epoch[0])
        elif gather == "p2p-nccl":
            scene.push_tiles(params_async, slab_t.data_ptr(), peer_ptrs)
            dist.all_reduce(token)
        else:
            dist.all_gather_into_tensor(gathered, slab_t)
            scene.untile(params_async, gathered.data_ptr(), canvas_dev.data_ptr())

    def frame_stats():
        st = scene.finish()
        return st, st["kernel_launches"] + (1 if world > 1 and not fused else 0)

    # N > 1, end to end: ONE host canvas in shared memory, page-locked in every rank's process. The image is dealt in
    # full-width bands of --band-h rows; each rank renders its bands and copies them straight into the shared canvas
    # over its own GPU's PCIe link (RT_RENDER_HOST_BANDS) — no device-side gather on this path.
    shared, shared_handle = None, None
    band_w = (W + 7) // 8 * 8
    if world > 1:
        shm_name = "/rtb200_bench_%s" % os.environ.get("MASTER_PORT", "0")
        if rank == 0:
            shared, shared_handle = pkg.host_canvas_open(W * H, shm_name=shm_name, create=True)
        dist.barrier()
        if rank != 0:
            shared, shared_handle = pkg.host_canvas_open(W * H, shm_name=shm_name)
        dist.barrier()
    staged = pkg.RT_RENDER_HOST_DIRECT if a.host_direct else 0
    e2e_flags = base | pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_HOST_BANDS | staged

    def frame_e2e():
        """The call a user of the reference makes per frame, with host buffers on both sides."""
        scene.set_instances(desc.instances, desc.n_instances)
        scene.set_camera(desc.camera)
        if world == 1:
            _, st = scene.render(flags=base | staged, tile_w=a.tile_w, tile_h=a.tile_h, out_ptr=host_canvas.data_ptr(),
                                 host_chunks=a.host_chunks)
            return st
        _, st = scene.render(flags=e2e_flags, rank=rank, world=world, tile_w=band_w, tile_h=a.band_h,
                             out_ptr=shared.ctypes.data, host_chunks=a.host_chunks)
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, sampler=None):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        total_ms, out = 0.0, []
        barrier()
        if sampler:
            sampler.start()
        for _ in range(steps):
            flush.zero_()            # L2 flush between timed iterations (untimed)
            if world > 1:
                dist.barrier()       # ranks enter the step together: the all-gather must not time rank skew
            torch.cuda.synchronize()
            ev0.record()
            r = fn()
            ev1.record()
            ev1.synchronize()
            out.append(r if r is not None else frame_stats())
            total_ms += ev0.elapsed_time(ev1)
        clocks = sampler.stop() if sampler else None
        barrier()
        if world > 1:
            t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
        return total_ms, out, clocks

    if gather in ("p2p", "p2p-nccl") or fused:
        # untimed check: the peer-memory exchange leaves the frame the all-gather path leaves
        frame_device()
        torch.cuda.current_stream().synchronize()
        frame_stats()
        receives = not fused or gather == "fused-all" or rank == 0
        src_ptr = scene.peer_frame_canvas(rank)[0] if fused else own_canvas_ptr
        got = scene.copy_to_host(src_ptr, np.empty(H * W, np.uint32))
        scene.render(flags=dev_flags | pkg.RT_RENDER_ASYNC, rank=rank, world=world, tile_w=a.tile_w, tile_h=a.tile_h,
                     out_ptr=slab_t.data_ptr())
        dist.all_gather_into_tensor(gathered, slab_t)
        scene.untile(params_async, gathered.data_ptr(), canvas_dev.data_ptr())
        torch.cuda.current_stream().synchronize()
        frame_stats()
        want = canvas_dev.cpu().numpy().view(np.uint32)
        if receives and not (got == want).all():
            raise SystemExit("rank %d: peer-memory frame differs from the all-gather frame in %d pixels" % (rank, int((got != want).sum())))
    # untimed check of the end-to-end paths against the device-resident frame
    scene.render(flags=dev_flags & ~pkg.RT_RENDER_TILED_OUTPUT, tile_w=a.tile_w, tile_h=a.tile_h, out_ptr=canvas_dev.data_ptr())
    want = canvas_dev.cpu().numpy().view(np.uint32)
    frame_e2e()
    barrier()
    got = host_canvas.numpy().view(np.uint32) if world == 1 else shared
    if rank == 0 and not (got == want).all():
        raise SystemExit("end-to-end frame differs from the device-resident frame in %d pixels" % int((got != want).sum()))
    sampler = ClockSampler(local_rank) if rank == 0 and not os.environ.get("BENCH_NO_CLOCKS") else None
    if sampler:
        sampler.wait_ready()
    if world > 1:
        dist.barrier()
    for _ in range(max(a.warmup, 3)):
        frame_device()
        torch.cuda.current_stream().synchronize()
        frame_stats()
        frame_e2e()
    ms_dev, outs, clocks = timed(frame_device, a.steps, sampler)
    ms_e2e, outs_e2e, _ = timed(frame_e2e, a.steps)

    # rays: every rank counts its own tiles
    rays_local = sum(o[0]["rays_total"] for o in outs)
    launches = sum(o[1] for o in outs)
    if world > 1:
        t = torch.tensor([rays_local, launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        rays_all, launches_all = int(t[0].item()), int(t[1].item())
    else:
        rays_all, launches_all = rays_local, launches
    rays_per_frame = rays_all / a.steps
    value = rays_all / (ms_dev * 1e-3) / 1e6
    e2e_value = rays_all / (ms_e2e * 1e-3) / 1e6

    # one stats frame for nodes / triangle tests per ray (slower kernels, untimed). Every rank takes
    # part: the frame contains the all-gather.
    frame_device(dev_flags | pkg.RT_RENDER_STATS)
    torch.cuda.current_stream().synchronize()
    sst, _ = frame_stats()
    if rank == 0:
        st = [o[0] for o in outs]
        avg = lambda k: sum(s[k] for s in st) / len(st)
        peak, peak_src = hbm_peak()
        l2_peak = pkg.measure_l2_bandwidth(32 << 20, 40)
        # dominant kernel and its algorithmic HBM bytes per launch (DESIGN.md "Roofline"):
        #   k_primary: 4 B per missed pixel (canvas) + 32 B per hit (hit queue write)
        pix_local = sst["rays_primary"]
        hits = sst["n_hits"]
        #   k_shadow : 32 B per traced shadow ray (queue read) + 1 B (occlusion flag)
        shadow_traced = sst["rays_shadow"] - sst["rays_culled"]
        if sst["tile_kernel"] == 1:
            # k_tile keeps rays, hits and shadow rays in registers: what it must move through HBM is the pixel it stores
            kernels = {"k_tile": (avg("ms_primary"), 4.0 * pix_local)}
        else:
            kernels = {("k_tile<1>" if sst["tile_kernel"] == 2 else "k_primary"): (avg("ms_primary"), 4.0 * (pix_local - hits) + 32.0 * hits),
                       "k_shadow": (avg("ms_shadow"), 33.0 * shadow_traced)}
        dom = max(kernels, key=lambda k: kernels[k][0])
        dom_ms, dom_bytes = kernels[dom]
        achieved = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
        l2_bytes = 80.0 * sst["nodes_visited"] + 48.0 * sst["prims_prefiltered"]
        ms_trace = avg("ms_trace")
        traffic, traffic_src = None, None
        try:  # dram__bytes_read.sum + dram__bytes_write.sum of that kernel, one launch, from the committed ncu capture
            cap = json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")))
            if cap["workload"] == workload_name(a) and world == 1:
                traffic, traffic_src = cap["kernels"][dom]["traffic"], cap["source"]
        except Exception:
            pass
        roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": dom_ms,
                    "note": "scene + BVH8 (< 2 MB) are L1/L2-resident; the kernel is issue/latency-bound, not HBM-bound",
                    "l2": {"bytes_per_frame": l2_bytes, "achieved": l2_bytes / (ms_trace * 1e-3) / 1e9,
                           "peak": l2_peak, "unit": "GB/s", "frac": l2_bytes / (ms_trace * 1e-3) / 1e9 / l2_peak,
                           "peak_source": "rt_measure_l2_bandwidth (32 MiB resident read, measured in this run)",
                           "nodes_per_ray": sst["nodes_visited"] / max(sst["rays_total"], 1),
                           "tri_filters_per_ray": sst["prims_prefiltered"] / max(sst["rays_total"], 1),
                           "exact_tests_per_ray": sst["tris_tested"] / max(sst["rays_total"], 1),
                           "node_fetches_per_ray_after_warp_broadcast": sst["node_fetches"] / max(sst["rays_total"], 1),
                           "bytes": "80 B x node visits counted per active ray + 48 B x triangle filters (SURVEY 8d). With k_tile a node is "
                                    "fetched once per warp and tested against 4 ray-group beams, so the bytes a per-ray traversal would "
                                    "request are no longer requested: node_fetches_per_ray is what is actually loaded"}}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
                "ms_per_step": ms_dev / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic (procedural world, seed 42)",
                "config": {"workload": workload_name(a), "rays_per_frame": rays_per_frame,
                           "rays": "the reference's cast_ray calls for this frame; culled_frac of them (shadow rays of lights "
                                   "with zero Phong weight) are answered without a traversal",
                           "culled_frac": sst["rays_culled"] / max(sst["rays_total"], 1), "instances": desc.n_instances,
                           "triangles": sst_prims(scene), "step": "BVH rebuild + trace" if not a.keep_bvh else "trace (BVH kept)",
                           "l2": "flushed between timed iterations (256 MiB write)",
                           "lanes": a.lanes if a.lanes else "library default (one stream unless RT_FRAME_LANES is set)",
                           "sharding": "screen tiles %dx%d round-robin over %d rank(s)" % (a.tile_w, a.tile_h, world),
                           "exchange": {"fused-root": "RT_RENDER_PEER_STORE: the frame's kernels store every pixel into rank 0's row-major canvas over NVLink peer memory (double-buffered by frame parity) and signal its mailbox; rank 0 waits for all ranks (the framebuffer gathered: no slab, no push kernel, no collective call)",
                                        "fused-all": "RT_RENDER_PEER_STORE into every rank's canvas (all ranks receive the frame)",
                                        "p2p": "tiles stored into every rank's canvas over NVLink peer memory, completion signalled through mailboxes in the same peer memory (rt_push_tiles_signal: no collective call on the frame's path)",
                                        "p2p-nccl": "tiles stored into every rank's canvas over NVLink peer memory (rt_push_tiles) + 1-element all-reduce as barrier",
                                        "nccl": "NCCL all-gather of the tile slabs + de-interleave kernel", "none": "single GPU"}[gather]},
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / a.steps, "h2d_bytes_per_step": h2d * world,
                        "d2h_bytes_per_step": 4 * W * H,
                        "path": e2e_path(world, a, outs_e2e[0]["host_chunks"])},
                "gpu_launches": launches_all, "clocks": clocks,
                "graph_frames": {"replayed": sum(1 for x in st if x["graph"] == 2), "captured": sum(1 for x in st if x["graph"] == 1), "plain": sum(1 for x in st if x["graph"] == 0)},
                "breakdown_ms": {"build": avg("ms_build"), "trace": ms_trace, "k_primary": avg("ms_primary"),
                                 "shade": avg("ms_shade"), "k_shadow": avg("ms_shadow"), "bounce_waves": avg("ms_bounce"),
                                 "waves": st[0]["waves"]},
                "roofline": roofline}
        line["frame_kernels"] = {0: "k_primary (per-lane walk) + k_shadow_gen + k_shadow + k_resolve (round 1's split pipeline)",
                                 1: "k_tile<1> + k_tile<2> (tile kernels, warp-coherent traversal, no shadow queues)",
                                 2: "k_tile<1> (warp-coherent closest hits) + k_shadow_gen + k_shadow + k_resolve"}[sst["tile_kernel"]]
        if world == 1 and not a.no_ref_cuda:
            line["ref_cuda"] = ref_cuda_block(a, a.scene, rays_per_frame, ms_dev / a.steps, ms_e2e / a.steps)
        if world == 1 and not a.no_stress and a.scene != "world8_stress":
            line["stress"] = single_gpu_block(pkg, torch, a, "world8_stress", max(5, a.steps // 2), flush)
        if world == 1 and not a.no_fast_accept:
            # the opt-in tolerance mode beside the bit-exact default (never the headline): same frames with RT_RENDER_FAST_ACCEPT
            def frame_fast():
                scene.render(flags=dev_flags | pkg.RT_RENDER_FAST_ACCEPT | pkg.RT_RENDER_ASYNC, tile_w=a.tile_w, tile_h=a.tile_h, out_ptr=canvas_dev.data_ptr())
            exact_canvas = want
            for _ in range(3):
                frame_fast(); torch.cuda.current_stream().synchronize(); frame_stats()
            ms_fast, outs_fast, _ = timed(frame_fast, max(5, a.steps // 2))
            fast_canvas = canvas_dev.cpu().numpy().view(np.uint32)
            chn = lambda c: np.stack([(c >> sh) & 255 for sh in (24, 16, 8, 0)], -1).astype(np.float64) / 255.0
            nf = len(outs_fast)
            line["fast_accept"] = {"flag": "RT_RENDER_FAST_ACCEPT (opt-in tolerance mode: clear filter hits skip the reference-exact test; default stays bit-exact)",
                                   "ms_per_step": ms_fast / nf, "value": sum(o[0]["rays_total"] for o in outs_fast) / (ms_fast * 1e-3) / 1e6, "unit": UNIT,
                                   "rmse_vs_exact_frame": float(np.sqrt(np.mean((chn(fast_canvas) - chn(exact_canvas)) ** 2))),
                                   "pixels_differ": int((fast_canvas != exact_canvas).sum()), "pixels": int(W * H),
                                   "parity_table": "profiles/r2_fast_accept_parity.md (ids 100 %, t within 8e-7 relative, RMSE <= 1.3e-4 on the five BASELINE configurations)"}
        if world == 1 and not a.no_cpu_baseline:
            v, kind, cores, sample, _ = cpu_reference_run(a, a.cpu_seconds)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
                                    "host_cores": os.cpu_count()}
        print(json.dumps(line), flush=True)
    if shared_handle:
        dist.barrier()
        pkg.host_canvas_close(shared_handle, unlink=(rank == 0))
    if peer_ptrs:
        dist.barrier()
        sets = fused_sets if fused_sets else [peer_ptrs]
        if fused_sets:
            scene.set_peers([], 0, 0)
        for st_ in sets:
            for r, ptr in enumerate(st_):
                if r != rank:
                    pkg.peer_canvas_close(ptr)
        dist.barrier()
        for st_ in sets:
            pkg.peer_canvas_free(st_[rank])
    scene.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def e2e_path(world, a, chunks):
    where = "a pinned host canvas" if world == 1 else \
        "ONE pinned host canvas in shared memory (full-width bands of %d rows dealt to the ranks, RT_RENDER_HOST_BANDS)" % a.band_h
    if chunks == 0:
        return "rt_render: the kernels write their pixels straight into %s over PCIe while the frame renders; no device-side gather, no copy after the frame" % where
    return "rt_render into a device canvas, copied into %s in %d chunk(s) of tile rows (copy of chunk c overlaps the rendering of c + 1)" % (where, chunks)


def sst_prims(scene):
    try:
        return scene.build_bvh()["n_prims"]
    except Exception:
        return None


def main():
    a = parse_args()
    if a.impl == "reference":
        return run_reference_arm(a)
    return run_ours(a)


if __name__ == "__main__":
    sys.exit(main())
