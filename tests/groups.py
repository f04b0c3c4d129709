"""In-process stand-ins for the plumbing group of mayamatchmovesolver_b200.distributed
(one host thread per rank), used to exercise the N>1 host logic without N processes."""
import threading


class ThreadGroupShared:
    def __init__(self, world):
        self.world = world
        self.barrier = threading.Barrier(world)
        self.slots = [None] * world
        self.bcast = None


class ThreadGroup:
    def __init__(self, shared, rank):
        self.s, self.rank, self.world = shared, rank, shared.world

    def broadcast_bytes(self, b, src=0):
        if self.rank == src:
            self.s.bcast = b
        self.s.barrier.wait()
        out = self.s.bcast
        self.s.barrier.wait()
        return out

    def allgather(self, obj):
        self.s.slots[self.rank] = obj
        self.s.barrier.wait()
        out = list(self.s.slots)
        self.s.barrier.wait()
        return out


def run_ranks(world, fn):
    """fn(group) on `world` threads -> list of return values (exceptions re-raised)"""
    shared = ThreadGroupShared(world)
    out, err = [None] * world, [None] * world

    def body(r):
        try:
            out[r] = fn(ThreadGroup(shared, r))
        except BaseException as e:  # noqa: BLE001
            err[r] = e
            shared.barrier.abort()

    ts = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    for e in err:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    for e in err:
        if e is not None:
            raise e
    return out
