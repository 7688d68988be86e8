#!/usr/bin/env python3
"""BASELINE config 5 on N GPUs of one node: soup-10M (generator of SURVEY.md 8d), 3840x2160, 64 spp per GPU,
bounce limit 4, sun on. One process per GPU (torchrun); every rank builds the tree on its own GPU (scene replicated:
480 MB of positions + wide nodes), renders its sample range, ONE NCCL reduce of the accumulation buffer per frame.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 \
        tools/soup_multi.py [--tris 10000000] [--spp 64] [--steps 2]

Rank 0 prints one JSON line: aggregate Mrays/s (device events, max over ranks), build ms (max over ranks)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=10_000_000)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import dustraytracer_b200 as drt
    from bench import _DevPtr
    from run_configs import soup
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    t0 = time.perf_counter()
    scene = drt.Scene.from_triangles(soup(a.tris))
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    drt.BVHBuilder(device=local).buildIterative(scene)
    build_wall = time.perf_counter() - t0
    st = scene.stats()
    r = drt.Renderer(device=local)
    r.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=1 << 30)
    r.ResizeBuffer(a.width, a.height)
    stream = torch.cuda.current_stream(dev)
    r.set_stream(stream.cuda_stream)
    r.set_frame_schedule(rank + 1, world)
    ptr, nbytes = r.accum_device_ptr()
    acc_t = torch.as_tensor(_DevPtr(ptr, nbytes // 4), device=dev)
    cam = drt.Camera(pos=(0.0, 0.0, 120.0))

    def step():
        r.resetAccumulationBuffer()
        r.Render(cam, scene, a.spp)
        if world > 1:
            dist.reduce(acc_t, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                r.resolve(a.spp * world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    step()
    r.counters(reset=True)
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    for k in range(a.steps):
        ev[k][0].record(stream)
        step()
        ev[k][1].record(stream)
    barrier()
    ms = sum(x.elapsed_time(y) for x, y in ev)
    c = r.counters()
    v = torch.tensor([float(c["closest_rays"] + c["shadow_rays"]), ms, st["build_ms"], build_wall], dtype=torch.float64, device=dev)
    if world > 1:
        tot = v.clone()
        dist.all_reduce(tot[0:1], op=dist.ReduceOp.SUM)
        dist.all_reduce(tot[1:4], op=dist.ReduceOp.MAX)
        v = tot
    if rank == 0:
        rays, ms, build_ms, bw = [float(x) for x in v]
        print(json.dumps({"config": "5", "scene": "soup:%d" % a.tris, "n_gpus": world, "width": a.width, "height": a.height,
                          "spp_per_gpu": a.spp, "steps": a.steps, "Mrays_s": rays / ms / 1e3, "ms_per_step": ms / a.steps,
                          "Msamples_s": a.width * a.height * a.spp * world * a.steps / ms / 1e3,
                          "build_ms_max": build_ms, "build_wall_s_max": bw, "generate_s": gen_s,
                          "wide_nodes": st["wide_nodes"], "device_MB": st["device_bytes"] / 1e6,
                          "parallelism": "sample-range sharding x%d + NCCL reduce of the accumulation buffer" % world}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
