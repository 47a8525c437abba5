"""CUDA-graph capture of the per-frame launch sequences.

One coded P-frame is ~400 kernel launches with fixed shapes; issued one by one from Python they cost
20-30 ms of host time per frame, which bounds throughput once several GPUs share one host (8 ranks x
8 interleaved sequences on 16 physical cores).  The device work of a frame is therefore captured
ONCE per (shape, q, reference-state) into CUDA graphs and replayed:

  * the encoder side of a frame (`DMC._analyse`, video_model.py:263-331) has no host hand-off: one graph;
  * the decoder side (video_model.py:354-422) hands index planes to the host rANS decoder four times
    (common_model.py:182-217), so it is a chain of graphs with the host coder between them.

A frame body is written once, as a generator: it does device work through the backend, `yield`s a
request whenever it needs symbols from the host coder, is sent back a device tensor holding them,
and returns its result.  `drive_eager` runs such a body directly; `ProgramCache.run` runs it
eagerly the first time a key is seen (that pass packs weights, sizes persistent buffers and sets
kernel attributes -- none of which may happen inside a capture), captures it the second time
(every segment between two yields becomes one graph, all in one private memory pool), and replays
the captured chain from then on.

Requests:  ("plane", idx)            idx: uint8 device tensor; reply int16 tensor of the same shape
           ("z", table, C, (h, w))   reply int16 [1, C, h, w]
`service(req)` does the host work and returns the reply as a HOST tensor.

Captured programs read and write fixed addresses: inputs are copied into static buffers by the
caller (models.py `_statics`), results are the same tensors on every replay.  All programs of one
model share a pool, so one model instance must be driven from one stream at a time (as its
persistent scratch buffers already require).
"""
import threading

import torch

_CAPTURE_LOCK = threading.Lock()


def drive_eager(gen, service, device):
    reply = None
    try:
        while True:
            req = gen.send(reply)
            reply = service(req).to(device, non_blocking=True)
    except StopIteration as e:
        return e.value


def _reply_like(req, device):
    if req[0] == "plane":
        return torch.empty(tuple(req[1].shape), dtype=torch.int16, device=device)
    if req[0] == "z":
        return torch.empty((1, req[2], req[3][0], req[3][1]), dtype=torch.int16, device=device)
    raise ValueError("unknown request %r" % (req[0],))


class Program:
    """A captured frame body: [(graph, request after it or None, static reply buffer, launches)]."""

    def __init__(self, backend):
        self.B = backend
        self.segs = []
        self.result = None

    def replay(self, service):
        B = self.B
        for g, req, up, n in self.segs:
            g.replay()
            B.launches += n
            B.graph_replays += 1
            if req is not None:
                up.copy_(service(req), non_blocking=True)    # pinned host buffer (models._Staging)
        return self.result


class ProgramCache:
    def __init__(self, backend):
        self.B = backend
        self.pool = None
        self.side = None
        self.seen = set()
        self.progs = {}

    MAX_PROGRAMS = 32      # distinct (shape, q, reference-state) keys kept; beyond that start over

    def invalidate(self):
        """Drop every captured program (their graphs bake in the addresses of packed weights, q
        vectors and static buffers): called when the model's parameters are reloaded."""
        self.progs.clear()
        self.seen.clear()
        self.pool = None       # the pool dies with its last graph: captures after this get a fresh one

    def run(self, key, make_gen, service):
        prog = self.progs.get(key)
        if prog is None:
            if key not in self.seen:
                self.seen.add(key)
                return drive_eager(make_gen(), service, self.B.device)
            if len(self.progs) >= self.MAX_PROGRAMS:
                self.progs.clear()
                self.pool = None
            prog = self._capture(make_gen)
            self.progs[key] = prog
        return prog.replay(service)

    def _capture(self, make_gen):
        B = self.B
        dev = B.device
        cur = torch.cuda.current_stream(dev)
        if self.pool is None:
            self.pool = torch.cuda.graph_pool_handle()
        if self.side is None or self.side.priority != cur.priority:
            # kernel nodes inherit the capture stream's priority (pipeline.py runs the decoder chain,
            # the critical path of a single sequence, on a high-priority stream)
            self.side = torch.cuda.Stream(dev, priority=cur.priority)
        prog = Program(B)
        gen = make_gen()
        reply = None
        launches0 = B.launches
        with _CAPTURE_LOCK:
            self.side.wait_stream(cur)
            with torch.cuda.stream(self.side):
                while True:
                    g = torch.cuda.CUDAGraph()
                    n0 = B.launches
                    g.capture_begin(pool=self.pool, capture_error_mode="thread_local")
                    try:
                        try:
                            req = gen.send(reply)
                            done = False
                        except StopIteration as e:
                            prog.result, req, done = e.value, None, True
                    except BaseException:
                        try:
                            g.capture_end()
                        except Exception:   # noqa: BLE001 -- the original error is the useful one
                            pass
                        raise
                    g.capture_end()
                    reply = None if done else _reply_like(req, dev)
                    prog.segs.append((g, req, reply, B.launches - n0))
                    if done:
                        break
            cur.wait_stream(self.side)
        B.launches = launches0          # a capture pass launches nothing; replays count
        return prog
