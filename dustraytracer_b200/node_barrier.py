"""Host barrier for the processes of ONE node, over a POSIX shared-memory segment.

The shared-frame-buffer mode of image-tile sharding (drt_renderer_import_frame_buffers) needs no data collective: its only
cross-rank step is "every rank's launch has returned". A NCCL barrier costs a kernel launch and a device synchronisation per
rank (~50-100 us), several times the render time of a one-sample frame split over eight GPUs; this one is a few
microseconds: rank r publishes the epoch it has reached in its own cache line and spins until every rank has.
"""
import time
from multiprocessing import shared_memory

import numpy as np

_LINE = 16  # uint64 per rank slot = 128 bytes: one cache line pair per writer


class NodeBarrier:
    def __init__(self, name, rank, world, create=None, timeout_s=60.0):
        """`name`: any string all ranks agree on (e.g. "drt-<MASTER_PORT>"). Rank 0 creates the segment; the others attach,
        retrying until it exists."""
        self.rank, self.world, self.timeout_s = rank, world, timeout_s
        self.epoch = 0
        nbytes = world * _LINE * 8
        create = (rank == 0) if create is None else create
        if create:
            try:
                stale = shared_memory.SharedMemory(name=name)
                stale.close()
                stale.unlink()
            except FileNotFoundError:
                pass
            self.shm = shared_memory.SharedMemory(name=name, create=True, size=nbytes)
            self.shm.buf[:nbytes] = bytes(nbytes)
        else:
            t0 = time.monotonic()
            while True:
                try:
                    self.shm = shared_memory.SharedMemory(name=name)
                    if self.shm.size >= nbytes:
                        break
                    self.shm.close()
                except FileNotFoundError:
                    pass
                if time.monotonic() - t0 > timeout_s:
                    raise TimeoutError("NodeBarrier: segment %r did not appear" % name)
                time.sleep(0.001)
            try:  # the creator unlinks the segment; keep this process's resource tracker from trying (and warning) at exit
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:  # noqa: BLE001
                pass
        self.owner = create
        self.slots = np.ndarray((world, _LINE), dtype=np.uint64, buffer=self.shm.buf)

    def wait(self):
        self.epoch += 1
        self.slots[self.rank, 0] = self.epoch
        col = self.slots[:, 0]
        t0 = None
        spins = 0
        while int(col.min()) < self.epoch:
            spins += 1
            if spins & 0x3fff == 0:
                t0 = t0 or time.monotonic()
                if time.monotonic() - t0 > self.timeout_s:
                    raise TimeoutError("NodeBarrier: rank %d waited %.0f s at epoch %d (slots %s)" % (self.rank, self.timeout_s, self.epoch, col.tolist()))

    def close(self):
        self.slots = None
        try:
            self.shm.close()
            if self.owner:
                self.shm.unlink()
        except (FileNotFoundError, BufferError):
            pass
