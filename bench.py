#!/usr/bin/env python3
"""bench.py -- headline benchmark of the path-tracing hot path (BASELINE.json: Mrays/s & samples/s, room.glb 1080p).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, NCCL)

One STEP = one pass of the hot path over one batch: the BASELINE config-3 frame -- room.glb, 1920x1080,
256 samples per pixel, ray_bounce_limit 8, sun shadow rays on -- rendered from a reset accumulation buffer
by ONE batched launch of the render kernel through the C ABI (drt_renderer_render).
  value : Mrays/s (TraceRay + RayTest calls / device time, CUDA events on the launching stream), scene, camera
          and frame buffers resident in HBM. Multi-GPU: sample-range sharding, every rank renders its own 256
          frame indices of the same image (weak scaling: 256*N spp per step) followed by one NCCL reduce of the
          24.9 MB accumulation buffer to rank 0 inside the timed region.
  e2e   : the same step through the public API with HOST buffers: camera/settings structs from host memory
          in, the resolved RGBA32F frame read back to pinned host memory, all inside the timed region.
  --impl reference : the reference's own traversal/shade sources compiled host-side (oracle/_ref, or the C
          port when that library is absent), all host threads, same scene/camera/settings; each step is a
          bounded sample of the workload (1 of the 256 samples per pixel).
The oracle is used only for the cpu_baseline / reference legs; the timed product path is CUDA only.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOAD = {
    "scene": "room.glb", "width": 1920, "height": 1080, "spp": 256, "bounces": 8, "sun": True,
}
if os.environ.get("DRT_BENCH_SPP"):  # profiling runs under ncu only: a shorter launch of the same kernel
    WORKLOAD["spp"] = int(os.environ["DRT_BENCH_SPP"])
CPU_SAMPLE_FRAMES = 128   # cpu_baseline leg: ~10-20 s on the GPU box's host cores
REF_STEP_FRAMES = 32      # --impl reference: samples per pixel rendered per step (bounded sample of the 256)
SCENE_PATH = os.path.join(ROOT, "tests", "golden", "scenes", WORKLOAD["scene"])


def log(*a):
    print(*a, file=sys.stderr, flush=True)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); power.append(float(r[3]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "power_w_max": max(power), "samples": len(sm)}


_JSON_FD = None


def claim_stdout():
    """The contract is ONE JSON line on stdout. Native libraries loaded by this process print on their own (the
    reference's loader and destructors: Scene.cu:43-54), and C stdio may flush late, so file descriptor 1 is pointed
    at stderr for the whole run and the JSON line is written to a private duplicate of the original stdout."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit_json_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


class stdout_to_stderr:
    """The reference's loader prints to stdout (Scene.cu:43-54); the bench's stdout carries exactly one JSON line."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        try:
            ctypes.CDLL(None).fflush(None)
        except Exception:
            pass
        os.dup2(self.saved, 1)
        os.close(self.saved)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


def oracle_scene():
    """(oracle, scene handle) of the CPU checker for the bench workload: the host-compiled reference when its
    prebuilt library travelled with the snapshot, else the C port fed by the product's HOST loader."""
    import numpy as np
    import oracles as O
    orc = O.best_oracle()
    if orc.kind == "reference":
        with stdout_to_stderr():
            return orc, orc.load(SCENE_PATH, bins=8, leaf=20, iterative=False)
    import dustraytracer_b200 as drt
    s = drt.Scene()
    s.loadGLTFmodel(SCENE_PATH)
    mats = np.frombuffer(b"".join(bytes(m) for m in s.materials()), np.uint32).reshape(-1, 11).copy()
    return orc, orc.create(s.triangles(), mats, [s.texture(i) for i in range(s.stats()["textures"])])


def cpu_reference_step(orc, oh, frame, frames=1, threads=0):
    """One bounded sample of the workload on the host cores: `frames` of the 256 samples per pixel."""
    import oracles as O
    cam = O.default_camera_bytes()
    st = O.settings_bytes(sun=WORKLOAD["sun"], bounces=WORKLOAD["bounces"], max_samples=100000)
    _, rays, sec = orc.render(oh, cam, st, WORKLOAD["width"], WORKLOAD["height"], frame, frames, nthreads=threads)
    return rays[0] + rays[1], sec


def ref_cuda_baseline(frames=8):
    """The reference's own CUDA path (unmodified sources, nvcc sm_100a, oracle/_ref/libdrt_refcuda*.so) on the same
    GPU, workload and launch shape as its Renderer (one 8x8-block launch + event pair per sample). Reported beside
    our number (north_star: >= 10x the reference's CUDA traversal on one B200); None when the libraries did not travel."""
    import numpy as np
    import oracles as O
    out = {}
    for key, lib in (("O3", "libdrt_refcuda.so"), ("O3_fast_math", "libdrt_refcuda_fast.so")):
        path = os.path.join(ROOT, "oracle", "_ref", lib)
        if not os.path.exists(path):
            continue
        L = ctypes.CDLL(path)
        L.refcuda_load.restype = ctypes.c_void_p
        L.refcuda_render.restype = ctypes.c_double
        with stdout_to_stderr():
            h = L.refcuda_load(SCENE_PATH.encode(), 8, 20)  # editor values: 8 bins, 20-triangle leaves (EditorLayer.cpp:53-54)
            out[key] = _ref_cuda_run(L, h, frames)
            L.refcuda_free(ctypes.c_void_p(h))  # the reference's Scene destructor prints too
    return out or None


def _ref_cuda_run(L, h, frames):
    import numpy as np
    import oracles as O
    cam = O.default_camera_bytes()
    st = O.settings_bytes(sun=WORKLOAD["sun"], bounces=WORKLOAD["bounces"], max_samples=100000)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    W, H = WORKLOAD["width"], WORKLOAD["height"]
    L.refcuda_render(ctypes.c_void_p(h), P(cam), P(st), W, H, 1, 2, None)  # warm-up
    acc = np.zeros(W * H * 3, np.float32)
    ms = L.refcuda_render(ctypes.c_void_p(h), P(cam), P(st), W, H, 1, frames, P(acc))
    return {"ms_per_sample": ms / frames, "samples_per_s": W * H * frames / (ms * 1e-3), "frames": frames,
            "mean_ldr": float(acc.mean() / frames)}


def run_reference(args, rank):
    if rank != 0:
        return 0
    orc, oh = oracle_scene()
    cores = os.cpu_count() or orc.max_threads()  # explicit: torchrun exports OMP_NUM_THREADS=1
    for w in range(args.warmup):
        cpu_reference_step(orc, oh, 1 + w * REF_STEP_FRAMES, frames=REF_STEP_FRAMES, threads=cores)
    rays_total, sec_total = 0, 0.0
    t0 = time.perf_counter()
    for k in range(args.steps):
        rays, sec = cpu_reference_step(orc, oh, 1 + (args.warmup + k) * REF_STEP_FRAMES, frames=REF_STEP_FRAMES, threads=cores)
        rays_total += rays
        sec_total += sec
    wall = time.perf_counter() - t0
    mrays = rays_total / sec_total / 1e6
    samples = WORKLOAD["width"] * WORKLOAD["height"] * args.steps * REF_STEP_FRAMES / sec_total
    sample_desc = "%d of the 256 samples per pixel of the 1920x1080 frame per step, OpenMP over rows" % REF_STEP_FRAMES
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic camera path over the bundled room scene (fixture re-muxed from the reference model)",
        "samples_per_s": samples,
        "config": {"workload": "room.glb 1920x1080, 8 bounces, sun shadow rays; CPU arm: %d spp per step" % REF_STEP_FRAMES, **WORKLOAD},
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": orc.kind, "sample": sample_desc},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": wall,
    }
    emit_json_line(line)
    return 0


class _DevPtr:
    """Exposes a raw device pointer to torch (zero-copy) through __cuda_array_interface__."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    import dustraytracer_b200 as drt

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W, H, SPP = WORKLOAD["width"], WORKLOAD["height"], WORKLOAD["spp"]

    scene = drt.Scene()
    scene.loadGLTFmodel(SCENE_PATH)
    drt.BVHBuilder(device=local_rank).buildIterative(scene)
    sstats = scene.stats()
    cam = drt.Camera()
    r = drt.Renderer(device=local_rank)
    r.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=WORKLOAD["bounces"], max_samples=1 << 30)
    r.set_tuning(int(os.environ.get("DRT_VARIANT", "-1")), int(os.environ.get("DRT_BLOCK", "0")))  # -1 = automatic (variant 4 at this sample count)
    r.ResizeBuffer(W, H)
    stream = torch.cuda.current_stream(dev)
    r.set_stream(stream.cuda_stream)
    p2p = args.shard == "p2p"
    tiles = args.shard in ("tiles", "p2p")
    if tiles:
        # image-tile sharding (strong scaling of one frame): rank g owns the 8x4 tiles g, g+world, ... of every frame
        r.set_tile_schedule(rank, world)
    else:
        # sample-range sharding: rank g owns frame indices g+1, g+1+world, ...
        r.set_frame_schedule(rank + 1, world)
    if p2p and world > 1:
        # image tiles WITHOUT a collective: every rank's kernel stores its tiles straight into rank 0's frame buffers
        # (CUDA IPC mapping + peer access over NVLink); the only cross-rank traffic besides those stores is the barrier
        box = [r.export_frame_buffers() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        if rank != 0:
            r.import_frame_buffers(box[0])
        from dustraytracer_b200.node_barrier import NodeBarrier
        node_barrier = NodeBarrier("drt-bench-%s" % os.environ.get("MASTER_PORT", "0"), rank, world)
        dist.barrier()
    ptr, nbytes = r.accum_device_ptr()
    acc_t = torch.as_tensor(_DevPtr(ptr, nbytes // 4), device=dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)  # > 126 MB L2
    host_frame = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    host_frame_np = host_frame.numpy()

    def step(e2e):
        """reset -> one batched 256-spp launch -> (N>1: NCCL reduce) -> (rank 0: resolve) [-> host readback]"""
        r.resetAccumulationBuffer()          # (p2p: rank 0 zeroes the shared buffers, the others only their sample counters)
        if p2p and world > 1:
            if rank == 0:
                stream.synchronize()         # the shared buffers are zeroed ...
            node_barrier.wait()              # ... before anybody stores into them (host barrier in shared memory, a few us)
        r.Render(cam, scene, SPP)            # (returns with this rank's tiles stored)
        if p2p and world > 1:
            node_barrier.wait()              # the frame (sums and display) is complete in rank 0's buffers
        elif world > 1:
            dist.reduce(acc_t, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                r.resolve(SPP if tiles else SPP * world)
        if e2e and rank == 0:
            r.readback(drt.DRT_RGBA32F, out=host_frame_np)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        step(False)
    r.counters(reset=True)
    launches0 = r.kernel_launches()

    # ---- device-resident timing: K steps, each bracketed by CUDA events on the launching stream ----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    t_wall0 = time.perf_counter()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for k in range(args.steps):
        flush.fill_(float(k))            # evict L2 between timed iterations (untimed)
        torch.cuda.synchronize(dev)
        ev[k][0].record(stream)
        step(False)
        ev[k][1].record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    launches = r.kernel_launches() - launches0
    ms = sum(a.elapsed_time(b) for a, b in ev)
    c = r.counters(reset=True)
    rays = torch.tensor([float(c["closest_rays"] + c["shadow_rays"]), ms], dtype=torch.float64, device=dev)
    if world > 1:
        tot = rays.clone()
        dist.all_reduce(tot[0:1], op=dist.ReduceOp.SUM)
        dist.all_reduce(tot[1:2], op=dist.ReduceOp.MAX)
        rays = tot
    total_rays, max_ms = float(rays[0]), float(rays[1])

    # ---- end to end through the public API with host buffers ----
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(True)
    barrier()
    e2e_s = time.perf_counter() - t0
    c2 = r.counters(reset=True)
    rays2 = torch.tensor([float(c2["closest_rays"] + c2["shadow_rays"])], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(rays2, op=dist.ReduceOp.SUM)
    e2e_rays = float(rays2[0])

    if rank == 0:
        # ---- traversal statistics for the algorithmic-bytes term (untimed, small sample) ----
        rs = drt.Renderer(device=local_rank, flags=drt.DRT_FLAG_STATS)
        rs.m_RendererSettings = r.m_RendererSettings
        rs.ResizeBuffer(W, H)
        rs.Render(cam, scene, 1)
        cs = rs.counters()
        srays = cs["closest_rays"] + cs["shadow_rays"]
        V, T = cs["node_visits"] / srays, cs["tri_tests"] / srays
        bytes_per_ray = 80.0 * V + 48.0 * T           # SURVEY.md 8d; megakernel: no ray/hit records in HBM
        fb_bytes = W * H * 48.0                        # per launch: accum read + write, display write (16 B each)
        peaks, peak_kind = measured_peaks()
        per_launch_ms = max_ms / args.steps
        rays_per_launch = total_rays / args.steps / world
        achieved = (bytes_per_ray * rays_per_launch + fb_bytes) / (per_launch_ms * 1e-3) / 1e9
        ncu = ncu_capture_summary()
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                    "traffic": ncu.get("traffic"), "traffic_note": ncu.get("traffic_note"), "issue": ncu.get("issue"),
                    "peak_kind": peak_kind, "kernel": "k_render<STRICT=0,STATS=0,VARIANT=%d> (phased, one launch per step)" % max(4, int(os.environ.get("DRT_VARIANT", "4"))),
                    "algorithmic_bytes_per_ray": bytes_per_ray, "node_visits_per_ray": V, "tri_tests_per_ray": T,
                    "note": "scene is cache resident (42 KB): the binding roofline is SM issue rate, see profiles/ and DESIGN.md"}
        # ---- CPU baseline: bounded sample on the host cores ----
        cpu = None
        if world == 1:  # reported baselines are measured at N=1 only (torchrun also pins OMP_NUM_THREADS=1)
            orc, oh = oracle_scene()
            ncores = os.cpu_count() or orc.max_threads()
            crays, csec = cpu_reference_step(orc, oh, 1, frames=CPU_SAMPLE_FRAMES, threads=ncores)
            cpu = {"value": crays / csec / 1e6, "unit": "Mrays/s", "cores": ncores, "kind": orc.kind,
                   "sample": "%d of the 256 samples per pixel (frame indices 1..%d) of the same 1920x1080 frame, %.1f s"
                             % (CPU_SAMPLE_FRAMES, CPU_SAMPLE_FRAMES, csec)}
        value = total_rays / (max_ms * 1e-3) / 1e6
        refcuda = None
        try:
            if world == 1:
                refcuda = ref_cuda_baseline()
        except Exception as e:  # measurement aid only
            refcuda = {"error": str(e)}
        if refcuda and "error" not in refcuda:
            rays_per_sample = total_rays / args.steps / world / SPP   # same workload: rays are a property of the frame
            for v in refcuda.values():
                v["Mrays_s_at_our_ray_count"] = rays_per_sample / (v["ms_per_sample"] * 1e-3) / 1e6
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": max_ms / args.steps, "higher_is_better": True, "scaling": "strong" if tiles else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic camera path over the bundled room scene (fixture re-muxed from the reference model)",
            "samples_per_s": W * H * SPP * (1 if tiles else world) * args.steps / (max_ms * 1e-3),
            "config": {"workload": "room.glb 1920x1080, %d spp per %s per step, 8 bounces, sun shadow rays (BASELINE config 3)" % (SPP, "frame (all GPUs together)" if tiles else "GPU"),
                       **WORKLOAD, "parallelism": ("image-tile sharding x%d (8x4 tiles, interleaved), peer stores into rank 0's frame buffers over NVLink, no collective (host barrier in shared memory)" if p2p
                                                   else "image-tile sharding x%d (8x4 tiles, interleaved) + NCCL reduce of the accumulation buffer" if tiles
                                                   else "sample-range sharding x%d + NCCL reduce of the accumulation buffer") % world,
                       "l2": "256 MB buffer written between timed steps (L2 flush)", "wide_nodes": sstats["wide_nodes"],
                       "build_ms": sstats["build_ms"]},
            "e2e": {"value": e2e_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 152, "d2h_bytes_per_step": W * H * 16},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "reference_cuda_on_this_gpu": refcuda,
            "rays_per_step": total_rays / args.steps, "wall_s_timed_region": t_wall,
        }
        emit_json_line(line)
    if world > 1:
        dist.barrier()
        if p2p:
            node_barrier.close()
        dist.destroy_process_group()
    return 0


def ncu_capture_summary():
    """DRAM bytes per launch and issue statistics of the render kernel from the committed `ncu --set full` capture
    (profiles/rNN*_render_phased_raw.txt, latest; the capture itself is never part of a timed run)."""
    import glob
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_render_phased_raw.txt")))
    if not found:
        return {}
    path = found[-1]  # the latest capture of the default kernel
    vals = {}
    for line in open(path):
        if " = " in line and not line.startswith("#"):
            k, v = line.rsplit(" = ", 1)
            name, _, unit = k.partition(" [")
            try:
                vals[name] = (float(v), unit.rstrip("]"))
            except ValueError:
                pass
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
    out = {}
    if "dram__bytes_read.sum" in vals and "dram__bytes_write.sum" in vals:
        out["traffic"] = sum(vals[k][0] * scale.get(vals[k][1], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        out["traffic_note"] = ("bytes per launch, dram read+write of one k_render launch under ncu --set full (16-spp launch; "
                               "a launch touches each frame-buffer pixel once whatever its sample count)")
    if "smsp__issue_active.avg.pct_of_peak_sustained_active" in vals:
        out["issue"] = {"issue_slot_fraction": vals["smsp__issue_active.avg.pct_of_peak_sustained_active"][0] / 100.0,
                        "active_lanes_per_instruction": vals.get("smsp__thread_inst_executed_per_inst_executed.ratio", (None,))[0],
                        "source": os.path.relpath(path, ROOT)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--shard", default=os.environ.get("DRT_SHARD", "samples"), choices=["samples", "tiles", "p2p"],
                    help="N>1: sample-range sharding (weak scaling, default), image-tile sharding of each frame + reduce (strong scaling), "
                         "or image tiles stored directly into rank 0's buffers over NVLink (p2p)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if not (args.impl != "reference" and world == 1 and args.gpus > 1):
        claim_stdout()  # (a torchrun re-launch below hands the real stdout to its rank 0 instead)
    if args.impl == "reference":
        return run_reference(args, rank)
    if world == 1 and args.gpus > 1:
        # convenience: re-launch under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
               "--master-port", os.environ.get("MASTER_PORT", "29511"), os.path.abspath(__file__), "--gpus", str(args.gpus),
               "--steps", str(args.steps), "--warmup", str(args.warmup), "--shard", args.shard]
        return subprocess.call(cmd)
    return run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
