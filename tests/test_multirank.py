"""The N>1 path on CPU: two `gloo` ranks run the sample-range sharding bench.py uses on the GPUs
(rank g renders frame indices g+1, g+1+world, ... into a private accumulation buffer; ONE reduce of the sums to
rank 0; display = sum / total samples). The per-pixel work is the device code compiled for the host
(tests/hostemu), so this checks the host-side sharding / reduce / resolve logic without a GPU."""
import os
import socket

import numpy as np
import pytest

import oracles as O

GOLDEN = os.path.join(O.ROOT, "tests", "golden")
W, H, FRAMES = 64, 36, 8


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, out_path):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = np.load(os.path.join(GOLDEN, "scene_room.npz"))
    emu = O.HostEmu()
    h = emu.load(O.scene_path("room"))
    per_rank = FRAMES // world
    # the frame schedule drt_renderer_set_frame_schedule(rank + 1, world) gives this rank
    acc, cnt = emu.render(h, g["camera"], g["settings"], W, H, rank + 1, world, per_rank, strict=True, rtl=True, steps=1)
    t = torch.from_numpy(acc.copy())
    rays = torch.tensor([float(cnt[0]), float(cnt[1])], dtype=torch.float64)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)          # the one collective of the path
    dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    if rank == 0:
        display = t.numpy() / np.float32(per_rank * world)  # drt_renderer_resolve(total samples)
        np.savez(out_path, acc=t.numpy(), display=display, rays=rays.numpy())
    dist.barrier()
    dist.destroy_process_group()
    emu.free(h)


@pytest.mark.parametrize("world", [2])
def test_two_gloo_ranks_shard_the_frames_and_reduce(tmp_path, world):
    import torch.multiprocessing as mp
    out_path = str(tmp_path / "rank0.npz")
    mp.start_processes(_rank_main, args=(world, _free_port(), out_path), nprocs=world, join=True, start_method="spawn")
    got = np.load(out_path)
    g = np.load(os.path.join(GOLDEN, "scene_room.npz"))
    emu = O.HostEmu()
    h = emu.load(O.scene_path("room"))
    full, cnt = emu.render(h, g["camera"], g["settings"], W, H, 1, 1, FRAMES, strict=True, rtl=True, steps=1)
    # same set of samples; only the fp32 summation order differs (shard by shard instead of frame by frame)
    tol = FRAMES * np.finfo(np.float32).eps * np.abs(full).max()
    assert np.abs(got["acc"] - full).max() <= tol
    assert int(got["rays"][0]) == cnt[0] and int(got["rays"][1]) == cnt[1]
    assert np.allclose(got["display"], full / np.float32(FRAMES), atol=tol)
    # and each shard is exactly the frames it owns, rendered alone
    for rank in range(world):
        alone = np.zeros(W * H * 3, np.float32)
        for f in range(rank + 1, FRAMES + 1, world):
            emu.render(h, g["camera"], g["settings"], W, H, f, 1, 1, strict=True, rtl=True, steps=1, acc=alone)
        shard, _ = emu.render(h, g["camera"], g["settings"], W, H, rank + 1, world, FRAMES // world, strict=True, rtl=True, steps=1)
        assert np.array_equal(alone.view(np.uint32), shard.view(np.uint32))
    emu.free(h)


def _tile_rank_main(rank, world, port, out_path):
    import torch
    import torch.distributed as dist
    import sys
    sys.path.insert(0, O.ROOT)
    from dustraytracer_b200.api import tile_owner_mask
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = np.load(os.path.join(GOLDEN, "scene_room.npz"))
    emu = O.HostEmu()
    h = emu.load(O.scene_path("room"))
    # drt_renderer_set_tile_schedule(rank, world): every rank renders ALL frames of the tiles it owns; the pixels of the
    # other ranks stay zero. (The host emulation renders the whole image; the device kernels skip the other tiles.)
    acc, cnt = emu.render(h, g["camera"], g["settings"], W, H, 1, 1, FRAMES, strict=True, rtl=True, steps=1)
    own = tile_owner_mask(W, H, rank, world)
    acc = np.where(np.repeat(own.reshape(-1), 3), acc, np.float32(0))
    t = torch.from_numpy(acc.copy())
    pixels = torch.tensor([float(own.sum())], dtype=torch.float64)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)          # the one collective of the path
    dist.all_reduce(pixels, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.savez(out_path, acc=t.numpy(), pixels=pixels.numpy())
    dist.barrier()
    dist.destroy_process_group()
    emu.free(h)


@pytest.mark.parametrize("world", [2])
def test_two_gloo_ranks_shard_the_tiles_and_reduce(tmp_path, world):
    """Image-tile sharding: the tiles of the ranks partition the image and the reduced sums are the single-rank BITS."""
    import torch.multiprocessing as mp
    out_path = str(tmp_path / "rank0.npz")
    mp.start_processes(_tile_rank_main, args=(world, _free_port(), out_path), nprocs=world, join=True, start_method="spawn")
    got = np.load(out_path)
    g = np.load(os.path.join(GOLDEN, "scene_room.npz"))
    emu = O.HostEmu()
    h = emu.load(O.scene_path("room"))
    full, _ = emu.render(h, g["camera"], g["settings"], W, H, 1, 1, FRAMES, strict=True, rtl=True, steps=1)
    emu.free(h)
    assert int(got["pixels"][0]) == W * H
    assert np.array_equal(got["acc"].view(np.uint32), full.view(np.uint32))


def _barrier_rank(rank, world, name, rounds, out_dir):
    import sys
    sys.path.insert(0, O.ROOT)
    from dustraytracer_b200.node_barrier import NodeBarrier
    b = NodeBarrier(name, rank, world)
    log = []
    for k in range(rounds):
        # a rank may only see round k of the others once they have all entered round k
        with open(os.path.join(out_dir, "r%d" % rank), "w") as f:
            f.write(str(k))
        b.wait()
        seen = [int(open(os.path.join(out_dir, "r%d" % r)).read() or -1) for r in range(world)]
        log.append(min(seen) >= k)
        b.wait()
    b.close()
    assert all(log), log


def test_node_barrier_orders_three_processes(tmp_path):
    """The shared-memory barrier the shared-frame-buffer mode uses instead of a collective."""
    import torch.multiprocessing as mp
    world = 3
    mp.start_processes(_barrier_rank, args=(world, "drt-test-%d" % os.getpid(), 200, str(tmp_path)), nprocs=world, join=True, start_method="spawn")
