#!/usr/bin/env python3
"""bench.py -- headline benchmark of the path-tracing hot path (BASELINE.json: Mrays/s & samples/s, room.glb 1080p).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, NCCL)

One STEP = one pass of the hot path over one batch: the BASELINE config-3 frame -- room.glb, 1920x1080,
256 samples per pixel, ray_bounce_limit 8, sun shadow rays on -- rendered from a reset accumulation buffer
by ONE batched launch of the render kernel through the C ABI (drt_renderer_render).
  value : Mrays/s (TraceRay + RayTest calls / device time, CUDA events on the launching stream), scene, camera
          and frame buffers resident in HBM. Multi-GPU (default, strong scaling): ONE 256-spp frame per step, rank g
          renders the frame indices g+1, g+1+N, ... (256/N samples per pixel), then one NCCL reduce of the 33 MB
          accumulation buffer to rank 0 and the resolve, all inside the timed region. --scaling weak: 256 spp per
          GPU per step. --shard tiles | p2p | fused: image tiles + reduce, image tiles stored into rank 0's buffers
          over NVLink, sample ranges added into rank 0's buffer by the render kernels (no collective).
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
CPU_SAMPLE_FRAMES = 128   # cpu_baseline leg: at most this many of the 256 samples per pixel (~10 s on 16 host cores) ...
CPU_SAMPLE_BUDGET_S = float(os.environ.get("DRT_CPU_BUDGET_S", "30"))   # ... and at most this much wall time
REF_STEP_FRAMES = 1       # --impl reference: ONE of the 256 samples per pixel per step (8.1 M rays, ~70 ms on 16 cores)
REF_BUDGET_S = float(os.environ.get("DRT_REF_BUDGET_S", "45"))          # wall budget of the whole reference arm
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


def host_threads():
    """Threads the CPU arm may use: the affinity mask of this process (not os.cpu_count(), which ignores taskset / cgroup
    cpusets), further limited by a cgroup CPU quota when one is set."""
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    try:  # cgroup v2 quota: "max 100000" or "<quota> <period>"
        q, per = open("/sys/fs/cgroup/cpu.max").read().split()[:2]
        if q != "max":
            n = max(1, min(n, int(float(q) / float(per) + 0.5)))
    except Exception:
        pass
    return max(1, n)


def cpu_child_main(argv):
    """Child process of the CPU arms (`bench.py --_cpu-child frames_per_step steps warmup threads acc_path|-`): renders
    `frames_per_step` of the 256 samples per pixel of the workload per step with the CPU checker and prints one JSON line per
    finished step on stdout. Runs in its own process so that the parent can enforce a wall budget whatever the host does."""
    import numpy as np
    import oracles as O
    fps, steps, warmup, threads, acc_path = int(argv[0]), int(argv[1]), int(argv[2]), int(argv[3]), argv[4]
    out = os.dup(1)
    os.dup2(2, 1)  # the reference's loader prints to stdout (Scene.cu:43-54)

    def say(d):
        os.write(out, (json.dumps(d) + "\n").encode())

    orc, oh = oracle_scene()
    cam = O.default_camera_bytes()
    st = O.settings_bytes(sun=WORKLOAD["sun"], bounces=WORKLOAD["bounces"], max_samples=100000)
    W, H = WORKLOAD["width"], WORKLOAD["height"]
    acc = np.zeros(W * H * 3, np.float32)
    say({"ready": True, "kind": orc.kind, "threads": threads, "omp_max_threads": orc.max_threads()})
    frame = 1
    for k in range(-warmup, steps):
        if k == 0:
            acc[:] = 0.0
            frame = 1  # the timed steps are frame indices 1, 2, ... (what the GPU arm's first frames are)
        _, rays, sec = orc.render(oh, cam, st, W, H, frame, fps, acc=acc, nthreads=threads)
        frame += fps
        say({"step": k, "rays": int(rays[0] + rays[1]), "sec": sec, "frames": fps})
        if k >= 0 and acc_path != "-" and (k == steps - 1 or (k + 1) % 16 == 0):
            np.save(acc_path + ".tmp.npy", acc)
            os.replace(acc_path + ".tmp.npy", acc_path)
            say({"saved": k + 1})
    return 0


def run_cpu_arm(frames_per_step, steps, warmup, budget_s, threads, acc_path=None, tag="cpu"):
    """Runs the CPU checker in a child process under a wall budget; returns what completed:
    {kind, threads, steps_done, rays, sec, frames, saved_frames, note}. Never raises, never outlives budget_s + 5 s."""
    import select
    import signal
    res = {"kind": None, "threads": threads, "steps_done": 0, "rays": 0, "sec": 0.0, "frames": 0, "saved_frames": 0, "note": ""}
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = str(threads)     # torchrun exports OMP_NUM_THREADS=1
    env.setdefault("OMP_WAIT_POLICY", "passive")  # idle team members sleep: no spin-wait pathologies on a shared / throttled host
    env.pop("OMP_PROC_BIND", None)
    cmd = [sys.executable, os.path.abspath(__file__), "--_cpu-child", str(frames_per_step), str(steps), str(warmup), str(threads), acc_path or "-"]
    t0 = time.perf_counter()
    try:
        child = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=None, env=env, start_new_session=True)
    except Exception as e:
        res["note"] = "could not start the CPU child: %s" % e
        return res
    old_term = None

    def kill_child(*_):
        try:
            os.killpg(child.pid, signal.SIGKILL)  # exactly the process group started above
        except Exception:
            pass

    def on_term(signum, frm):
        kill_child()
        raise KeyboardInterrupt("signal %d" % signum)

    try:
        old_term = signal.signal(signal.SIGTERM, on_term)
    except Exception:
        pass
    buf = b""
    try:
        while True:
            left = budget_s - (time.perf_counter() - t0)
            if left <= 0:
                res["note"] = "wall budget of %.0f s reached after %d of %d steps" % (budget_s, res["steps_done"], steps)
                break
            rd, _, _ = select.select([child.stdout], [], [], min(left, 5.0))
            if not rd:
                log("[%s] waiting for the CPU child (%.0f s elapsed, %d steps done, threads %d)" % (tag, time.perf_counter() - t0, res["steps_done"], threads))
                continue
            chunk = os.read(child.stdout.fileno(), 65536)
            if not chunk:
                break
            buf += chunk
            while b"\n" in buf:
                ln, buf = buf.split(b"\n", 1)
                try:
                    d = json.loads(ln)
                except Exception:
                    continue
                if "ready" in d:
                    res["kind"] = d["kind"]
                    log("[%s] CPU child ready after %.1f s: kind %s, threads %d (affinity %d, os.cpu_count %s, omp_get_max_threads %d)"
                        % (tag, time.perf_counter() - t0, d["kind"], threads, len(os.sched_getaffinity(0)), os.cpu_count(), d["omp_max_threads"]))
                elif "saved" in d:
                    res["saved_frames"] = d["saved"] * frames_per_step
                elif "step" in d:
                    log("[%s] step %d: %.3f s, %.1f Mrays/s, threads %d, %.1f s elapsed" % (tag, d["step"], d["sec"], d["rays"] / max(d["sec"], 1e-9) / 1e6, threads, time.perf_counter() - t0))
                    if d["step"] >= 0:
                        res["steps_done"] += 1
                        res["rays"] += d["rays"]
                        res["sec"] += d["sec"]
                        res["frames"] += d["frames"]
            if res["steps_done"] >= steps and (acc_path is None or res["saved_frames"] >= res["frames"]):
                break
    except KeyboardInterrupt as e:
        res["note"] = "interrupted (%s) after %d of %d steps" % (e, res["steps_done"], steps)
    finally:
        kill_child()
        try:
            child.wait(timeout=5)
        except Exception:
            pass
        if old_term is not None:
            try:
                signal.signal(signal.SIGTERM, old_term)
            except Exception:
                pass
    return res


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
    """--impl reference: the reference's own traversal / shade sources compiled host-side (oracle/_ref; the C port when that
    library did not travel), all host threads of the affinity mask, ONE of the 256 samples per pixel per step, under a wall
    budget (REF_BUDGET_S). The JSON line is emitted from a `finally`, from whatever steps completed."""
    if rank != 0:
        return 0
    threads = host_threads()
    t0 = time.perf_counter()
    res = {"kind": None, "threads": threads, "steps_done": 0, "rays": 0, "sec": 0.0, "frames": 0, "note": "not started"}
    attempts = []
    try:
        warm = min(args.warmup, 16)  # W untimed steps as asked (bounded: a CPU step is ~70 ms on 16 threads, 1 s on one)
        res = run_cpu_arm(REF_STEP_FRAMES, args.steps, warm, REF_BUDGET_S, threads, tag="reference")
        attempts.append({"threads": threads, "steps_done": res["steps_done"], "note": res["note"]})
        if res["steps_done"] == 0 and threads > 1:
            # nothing finished with the full team: one thread, no OpenMP team at all
            log("[reference] no step completed with %d threads (%s); retrying single-threaded" % (threads, res["note"]))
            res = run_cpu_arm(REF_STEP_FRAMES, args.steps, 0, REF_BUDGET_S, 1, tag="reference/1-thread")
            attempts.append({"threads": 1, "steps_done": res["steps_done"], "note": res["note"]})
    finally:
        done = res["steps_done"]
        sec = res["sec"] if done else float("nan")
        mrays = res["rays"] / res["sec"] / 1e6 if done else 0.0
        W, H = WORKLOAD["width"], WORKLOAD["height"]
        sample_desc = ("%d of the 256 samples per pixel of the 1920x1080 frame per step (%.1f M rays), OpenMP over rows, %d threads; %d of %d steps completed%s"
                       % (REF_STEP_FRAMES, res["rays"] / max(done, 1) / 1e6, res["threads"], done, args.steps, ("; " + res["note"]) if res["note"] else ""))
        line = {
            "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": done,
            "steps_requested": args.steps, "warmup": min(args.warmup, 16), "ms_per_step": 1e3 * sec / max(done, 1), "higher_is_better": True,
            "scaling": "strong" if args.scaling == "strong" else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic camera path over the bundled room scene (fixture re-muxed from the reference model)",
            "samples_per_s": W * H * res["frames"] / res["sec"] if done else 0.0,
            "config": {"workload": "room.glb 1920x1080, 8 bounces, sun shadow rays (BASELINE config 3); CPU arm: %d spp per step, wall budget %.0f s" % (REF_STEP_FRAMES, REF_BUDGET_S), **WORKLOAD},
            "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": res["threads"], "kind": res["kind"] or "none", "sample": sample_desc},
            "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0, "attempts": attempts,
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
    strong = args.scaling == "strong" or args.shard in ("tiles", "p2p", "fused")
    if strong and args.shard in ("samples", "fused") and SPP % world:
        raise SystemExit("bench.py: strong sample-range sharding needs spp (%d) divisible by the number of GPUs (%d)" % (SPP, world))
    # samples per pixel THIS rank renders per step: strong scaling = one SPP-sample frame split over the ranks
    rank_spp = SPP // world if (strong and args.shard in ("samples", "fused")) else SPP
    frame_spp = SPP if strong else SPP * world  # samples per pixel of the frame one step produces (all GPUs together)

    scene = drt.Scene()
    scene.loadGLTFmodel(SCENE_PATH)
    drt.BVHBuilder(device=local_rank).buildIterative(scene)
    build_ms_cold = scene.stats()["build_ms"]   # first CUDA work of the process: context, module load, allocator
    drt.BVHBuilder(device=local_rank).buildIterative(scene)
    sstats = scene.stats()                       # the same build again, warm
    cam = drt.Camera()
    r = drt.Renderer(device=local_rank)
    r.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=WORKLOAD["bounces"], max_samples=1 << 30)
    r.set_tuning(int(os.environ.get("DRT_VARIANT", "-1")), int(os.environ.get("DRT_BLOCK", "0")))  # -1 = automatic (variant 4 at this sample count)
    r.ResizeBuffer(W, H)
    stream = torch.cuda.current_stream(dev)
    r.set_stream(stream.cuda_stream)
    fused = args.shard == "fused" and world > 1   # sample ranges, every rank's kernel ADDS its sums into rank 0's buffer (no collective)
    p2p = args.shard == "p2p" or fused           # frame buffers of rank 0 shared over CUDA IPC / NVLink
    tiles = args.shard in ("tiles", "p2p")
    if fused:
        r.set_accumulate_add(True)
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
        """reset -> one batched launch of this rank's share -> (N>1: NCCL reduce) -> (rank 0: resolve) [-> host readback]"""
        r.resetAccumulationBuffer()          # (p2p: rank 0 zeroes the shared buffers, the others only their sample counters)
        if p2p and world > 1:
            if rank == 0:
                stream.synchronize()         # the shared buffers are zeroed ...
            node_barrier.wait()              # ... before anybody stores into them (host barrier in shared memory, a few us)
        r.Render(cam, scene, rank_spp)       # (returns with this rank's share done: its tiles stored / its frame indices summed)
        if p2p and world > 1:
            node_barrier.wait()              # the frame (sums and display) is complete in rank 0's buffers
            if fused and rank == 0:
                r.resolve(frame_spp)         # (fused: the kernels added the sums; the display buffer is derived here)
        elif world > 1:
            dist.reduce(acc_t, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                r.resolve(frame_spp)
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

    # ---- N > 1: where the step goes (untimed extra steps; CUDA events on the launching stream around each part) ----
    breakdown = None
    if world > 1 and not p2p:
        parts = []
        for _ in range(5):
            r.resetAccumulationBuffer()
            barrier()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record(stream)
            r.Render(cam, scene, rank_spp)
            e[1].record(stream)
            dist.reduce(acc_t, dst=0, op=dist.ReduceOp.SUM)
            e[2].record(stream)
            if rank == 0:
                r.resolve(frame_spp)
            e[3].record(stream)
            torch.cuda.synchronize(dev)
            parts.append([e[i].elapsed_time(e[i + 1]) for i in range(3)])
        r.counters(reset=True)
        best = torch.tensor([min(p[i] for p in parts) for i in range(3)], dtype=torch.float64, device=dev)
        worst = best.clone()
        dist.all_reduce(worst[0:1], op=dist.ReduceOp.MAX)   # the slowest rank's launch is what the reduce waits for
        if rank == 0:
            breakdown = {"render_ms_slowest_rank": round(float(worst[0]), 4), "render_ms_rank0": round(float(best[0]), 4),
                         "reduce_ms_rank0": round(float(best[1]), 4), "resolve_ms_rank0": round(float(best[2]), 4),
                         "reduce_bytes": int(acc_t.numel() * 4),
                         "note": "best of 5 untimed steps started from a barrier; reduce = ncclReduce of the W*H float4 sums to rank 0 as seen by rank 0 "
                                 "(includes waiting for the slowest rank's launch); ideal render = one-GPU step / N"}

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
        algorithmic_bw = {"achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "peak_kind": peak_kind,
                          "algorithmic_bytes_per_ray": bytes_per_ray, "node_visits_per_ray": V, "tri_tests_per_ray": T,
                          "note": "SURVEY 8d byte formula (80 B x node visits + 48 B x triangle tests per ray + 48 B per pixel per launch); these bytes are served by L1 "
                                  "(scene = 42 KB), so this is NOT the binding roofline -- see `traffic` for what DRAM really moves"}
        # The binding roofline is SM issue rate x SIMT lanes: achieved = thread-level instructions per second =
        # (thread instructions per ray of the committed ncu capture of this very launch shape) x (rays per second measured here);
        # peak = 148 SMs x 4 schedulers x 32 lanes x the SM clock sampled during the timed region.
        sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0
        issue_peak = 148 * 4 * 32 * sm_mhz * 1e6
        rays_per_s = total_rays / (max_ms * 1e-3) / world   # per GPU
        tipr = None
        if ncu.get("thread_inst") and ncu.get("spp"):
            capture_rays = rays_per_launch * ncu["spp"] / rank_spp   # rays are a property of the frame: proportional to the sample count
            tipr = ncu["thread_inst"] / capture_rays
        issue_achieved = tipr * rays_per_s if tipr else None
        roofline = {"bound": "issue", "achieved": issue_achieved / 1e12 if issue_achieved else None, "peak": issue_peak / 1e12, "unit": "T thread-instructions/s",
                    "frac": issue_achieved / issue_peak if issue_achieved else None,
                    "traffic": ncu.get("traffic"), "traffic_note": ncu.get("traffic_note"),
                    "thread_instructions_per_ray": tipr, "warp_instructions_per_ray": (ncu["warp_inst"] / capture_rays) if tipr and ncu.get("warp_inst") else None,
                    "capture": {k: ncu.get(k) for k in ("source", "spp", "duration_ms", "issue_slot_fraction", "active_lanes_per_instruction", "local_sectors")},
                    "peak_note": "148 SMs x 4 schedulers x 32 lanes x %.0f MHz (SM clock sampled during the timed region)" % sm_mhz,
                    "kernel": "k_render<STRICT=0,STATS=0,VARIANT=%d> (per-warp phase voting, one launch per step; 22 = with the sample pool)" % r.last_variant(),
                    "algorithmic_bw": algorithmic_bw}
        # ---- CPU baseline: bounded sample on the host cores (child process under a wall budget), and the parity of the GPU's
        #      first frames of the SAME image against the frames the CPU just rendered ----
        cpu = None
        parity = None
        skip_baselines = bool(os.environ.get("DRT_BENCH_SKIP_BASELINES"))  # profiling runs under ncu
        if world == 1 and not skip_baselines:  # reported baselines are measured at N=1 only
            threads = host_threads()
            acc_path = os.path.join(ROOT, "gpurun_out", "cpu_baseline_acc_%d.npy" % os.getpid())
            os.makedirs(os.path.dirname(acc_path), exist_ok=True)
            # steps of 8 samples per pixel, up to CPU_SAMPLE_FRAMES samples, at most CPU_SAMPLE_BUDGET_S seconds
            res = run_cpu_arm(8, CPU_SAMPLE_FRAMES // 8, 0, CPU_SAMPLE_BUDGET_S, threads, acc_path=acc_path, tag="cpu_baseline")
            if res["steps_done"]:
                cpu = {"value": res["rays"] / res["sec"] / 1e6, "unit": "Mrays/s", "cores": res["threads"], "kind": res["kind"],
                       "sample": "%d of the 256 samples per pixel (frame indices 1..%d) of the same 1920x1080 frame, %.1f s%s"
                                 % (res["frames"], res["frames"], res["sec"], ("; " + res["note"]) if res["note"] else "")}
            else:
                cpu = {"value": None, "unit": "Mrays/s", "cores": threads, "kind": res["kind"], "sample": "no step completed: " + res["note"]}
            n_par = res["saved_frames"]
            if n_par and os.path.exists(acc_path):
                ref_acc = np.load(acc_path).reshape(H, W, 3).astype(np.float64) / n_par
                r.set_frame_schedule(1, 1)
                r.resetAccumulationBuffer()
                r.Render(cam, scene, n_par)            # production mode, the benchmarked kernel, frame indices 1..n_par
                got = r.readback(drt.DRT_ACCUM_RGB32F).reshape(H, W, 3).astype(np.float64) / n_par
                d = got - ref_acc
                parity = {"spp": n_par, "rmse": float(np.sqrt((d * d).mean())), "mean_diff": float(d.mean()), "max_abs_diff": float(np.abs(d).max()),
                          "pixels_differing": float((np.abs(d).max(axis=2) > 0).mean()),
                          "what": "display values (LDR, [0,1]) of the GPU's production-mode render of frame indices 1..%d of the benchmarked 1080p image "
                                  "against the CPU reference's render of the same frame indices (same seeds)" % n_par}
            try:
                os.remove(acc_path)
            except OSError:
                pass
        value = total_rays / (max_ms * 1e-3) / 1e6
        refcuda = None
        try:
            if world == 1 and not skip_baselines:
                refcuda = ref_cuda_baseline()
        except Exception as e:  # measurement aid only
            refcuda = {"error": str(e)}
        if refcuda and "error" not in refcuda:
            rays_per_sample = total_rays / args.steps / world / rank_spp   # same workload: rays are a property of the frame
            for v in refcuda.values():
                v["Mrays_s_at_our_ray_count"] = rays_per_sample / (v["ms_per_sample"] * 1e-3) / 1e6
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": max_ms / args.steps, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic camera path over the bundled room scene (fixture re-muxed from the reference model)",
            "samples_per_s": W * H * frame_spp * args.steps / (max_ms * 1e-3),
            "config": {"workload": "room.glb 1920x1080, %d spp per %s per step, 8 bounces, sun shadow rays (BASELINE config 3)" % (SPP, "frame (all GPUs together)" if strong else "GPU"),
                       **WORKLOAD, "parallelism": ("sample-range sharding x%d, fused render + reduce: every rank's kernel adds its pixel sums into rank 0's accumulation buffer with fp32 atomics over NVLink, no collective (host barrier in shared memory)" if fused
                                                   else "image-tile sharding x%d (8x4 tiles, interleaved), peer stores into rank 0's frame buffers over NVLink, no collective (host barrier in shared memory)" if p2p
                                                   else "image-tile sharding x%d (8x4 tiles, interleaved) + NCCL reduce of the accumulation buffer" if tiles
                                                   else "sample-range sharding x%d + NCCL reduce of the accumulation buffer") % world,
                       "spp_per_gpu_per_step": rank_spp, "spp_per_frame": frame_spp,
                       "l2": "256 MB buffer written between timed steps (L2 flush)", "wide_nodes": sstats["wide_nodes"],
                       "build_ms": sstats["build_ms"], "build_ms_cold": build_ms_cold, "build_launches": sstats.get("build_launches")},
            "e2e": {"value": e2e_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 152, "d2h_bytes_per_step": W * H * 16},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity": parity,
            "reference_cuda_on_this_gpu": refcuda,
            "rays_per_step": total_rays / args.steps, "wall_s_timed_region": t_wall,
        }
        if breakdown:
            line["step_breakdown"] = breakdown
        emit_json_line(line)
    if world > 1:
        dist.barrier()
        if p2p:
            node_barrier.close()
        dist.destroy_process_group()
    return 0


def ncu_capture_summary():
    """Counters of the render kernel from the latest committed ncu capture (profiles/rNN*_render_phased_raw.txt, written by
    tools/ncu_raw.py from a capture of `bench.py --steps 1`; its header says which launch: `spp=<n>`). The capture itself
    is never part of a timed run."""
    import glob
    import re
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_render_phased_raw.txt")))
    if not found:
        return {}
    path = found[-1]  # the latest capture of the default kernel
    vals = {}
    spp = None
    for line in open(path):
        if line.startswith("#"):
            m = re.search(r"spp=(\d+)", line) or re.search(r"DRT_BENCH_SPP=(\d+)", line)
            if m:
                spp = int(m.group(1))
        elif " = " in line:
            k, v = line.rsplit(" = ", 1)
            name, _, unit = k.partition(" [")
            try:
                vals[name] = (float(v), unit.rstrip("]"))
            except ValueError:
                pass
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
    out = {"source": os.path.relpath(path, ROOT), "spp": spp}
    if "dram__bytes_read.sum" in vals and "dram__bytes_write.sum" in vals:
        out["traffic"] = sum(vals[k][0] * scale.get(vals[k][1], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        out["traffic_note"] = ("bytes per launch, dram read+write of one k_render launch under ncu (%s-spp launch; "
                               "the frame buffer is touched once per pixel whatever the sample count; with the sample-pool kernel the rest "
                               "is the pool's colour scratch, 16 B per pooled sample written once and read back through L2)" % spp)
    get = lambda k: vals[k][0] if k in vals else None
    out["thread_inst"] = get("smsp__thread_inst_executed.sum")
    out["warp_inst"] = get("smsp__inst_executed.sum")
    out["duration_ms"] = get("gpu__time_duration.sum")
    if get("smsp__issue_active.avg.pct_of_peak_sustained_active") is not None:
        out["issue_slot_fraction"] = get("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0
    out["active_lanes_per_instruction"] = get("smsp__thread_inst_executed_per_inst_executed.ratio")
    if out["thread_inst"] is None and out["warp_inst"] and out["active_lanes_per_instruction"]:
        out["thread_inst"] = out["warp_inst"] * out["active_lanes_per_instruction"]
    ld, st = get("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum"), get("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum")
    if ld is not None and st is not None:
        out["local_sectors"] = ld + st
    return out


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--_cpu-child":
        return cpu_child_main(sys.argv[2:])
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--shard", default=os.environ.get("DRT_SHARD", "samples"), choices=["samples", "tiles", "p2p", "fused"],
                    help="N>1: sample-range sharding + NCCL reduce (default), image-tile sharding of each frame + reduce, image tiles stored "
                         "directly into rank 0's buffers over NVLink (p2p), or sample ranges added into rank 0's buffer by the render kernels (fused)")
    ap.add_argument("--scaling", default=os.environ.get("DRT_SCALING", "strong"), choices=["strong", "weak"],
                    help="N>1 with sample-range sharding: strong (default) = ONE 256-spp frame, 256/N samples per pixel per GPU + reduce + resolve; "
                         "weak = 256 spp per GPU (a 256*N-spp frame per step)")
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
               "--steps", str(args.steps), "--warmup", str(args.warmup), "--shard", args.shard, "--scaling", args.scaling]
        return subprocess.call(cmd)
    return run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
