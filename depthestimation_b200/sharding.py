"""Frame sharding across GPUs: frames are independent, so frame/chunk i goes to GPU
i mod G; every GPU has its own engine handle + stream + host worker thread, results are gathered on
the host in frame order. No collective on the data path.

Two drivers:
* ``ThreadedSharder``  — one process, G engines, G worker threads (what StereoDepthEstimatorVideo uses)
* ``frames_for_rank`` / ``gather_in_order`` — one process per GPU under torchrun (what bench.py uses);
  the gather is a host-side ``torch.distributed`` object gather (gloo is enough)
"""
from __future__ import annotations

import queue
import threading
from typing import Callable, Dict, Iterable, Iterator, List, Optional, Sequence, Tuple

import numpy as np


def frames_for_rank(n_frames: int, rank: int, world: int) -> List[int]:
    """Indices of the frames rank ``rank`` of ``world`` processes owns: i = rank (mod world)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("need 0 <= rank < world")
    return list(range(rank, n_frames, world))


def gather_in_order(local: Dict[int, object], n_frames: int, group=None, dst: int = 0) -> Optional[List[object]]:
    """Host-side gather of per-frame results to ``dst`` in frame order. Every rank passes the results of
    the frames it owns (index -> object). Returns the ordered list on ``dst``, None elsewhere.
    Raises if a frame is missing or produced twice."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    bucket = [None] * world if rank == dst else None
    dist.gather_object(local, bucket, dst=dst, group=group)
    if rank != dst:
        return None
    out: List[object] = [None] * n_frames
    seen = [False] * n_frames
    for part in bucket:
        for i, v in part.items():
            if not (0 <= i < n_frames) or seen[i]:
                raise RuntimeError(f"frame {i} gathered twice or out of range")
            out[i], seen[i] = v, True
    if not all(seen):
        raise RuntimeError(f"frames missing after gather: {[i for i, s in enumerate(seen) if not s][:8]}")
    return out


class ThreadedSharder:
    """Chunks of frames -> GPUs round-robin (chunk i to GPU i mod G), ordered results back.

    ``make_worker(device)`` is called once per device inside that device's worker thread (so each thread owns its
    handle, like one cv2.StereoSGBM per thread in the reference) and returns either
    * a callable ``fn(lefts[B,...], rights[B,...]) -> (disp[B,...] | None, depth[B,...] | None)`` (blocking), or
    * an object with ``submit(lefts: list, rights: list) -> token`` and ``wait(token) -> (disp, depth)`` and an
      optional ``max_inflight`` (default 3): the worker then keeps that many chunks in flight on its GPU, so uploads,
      kernels and downloads of consecutive chunks overlap (b200s_pipeline_frames_submit / b200s_pipeline_batch_wait).
      ``submit`` receives the chunk as two lists of frames (no host-side stacking on the dispatcher thread).
    A failing worker hands its chunks (in flight and waiting) to whichever of the others is free first, and later
    chunks that would have been its turn go the same way; when every worker has failed the error is raised to the caller.
    """

    def __init__(self, devices: Sequence[int], make_worker: Callable[[int], Callable], chunk: int = 8,
                 max_inflight: int = 4, ramp: bool = False):
        if not devices:
            raise ValueError("ThreadedSharder needs at least one device")
        self.devices, self.make_worker, self.chunk = list(devices), make_worker, max(1, int(chunk))
        self.max_inflight = max(1, int(max_inflight))
        # ramp: the first chunk of every GPU is a quarter, the second a half of `chunk`, so the GPUs start after a
        # quarter of the fill time of a full chunk (a 240-frame job is only three full chunks of 78)
        self.ramp = bool(ramp)

    def run(self, pairs: Iterable[Tuple[np.ndarray, np.ndarray]]) -> Iterator[Tuple[np.ndarray, Optional[np.ndarray]]]:
        G = len(self.devices)
        cv = threading.Condition()                 # guards everything below
        mine: List[List[tuple]] = [[] for _ in range(G)]   # chunk i waits in the list of worker i mod G (round-robin) ...
        orphans: List[tuple] = []                  # ... and the chunks of a failed worker here, where any worker takes them
        dead = [False] * G
        stop = [False]
        done: Dict[int, object] = {}
        errors: List[BaseException] = []

        def fail(w, exc, items):
            with cv:
                errors.append(exc)
                dead[w] = True
                orphans.extend(items)
                orphans.extend(mine[w])
                mine[w].clear()
                orphans.sort(key=lambda it_: it_[0])
                cv.notify_all()

        def take(w, block):
            """Next chunk for worker w (orphans first: they are the oldest), None when there is none (yet)."""
            with cv:
                while True:
                    if orphans:
                        return orphans.pop(0)
                    if mine[w]:
                        return mine[w].pop(0)
                    if stop[0] or not block:
                        return None
                    cv.wait(timeout=0.5)

        def worker(w, dev):
            try:
                fn = self.make_worker(dev)
            except BaseException as exc:  # noqa: BLE001
                fail(w, exc, [])
                return
            if not hasattr(fn, "submit"):        # blocking worker: one chunk at a time
                while True:
                    item = take(w, True)
                    if item is None:
                        break
                    cid, Ls, Rs = item
                    try:
                        res = fn(np.ascontiguousarray(np.stack(Ls)), np.ascontiguousarray(np.stack(Rs)))
                    except BaseException as exc:  # noqa: BLE001
                        fail(w, exc, [item])
                        return
                    with cv:
                        done[cid] = res
                        cv.notify_all()
                return
            cap = max(1, int(getattr(fn, "max_inflight", 3)))
            flying = []                          # [(item, token)] oldest first
            while True:
                try:
                    while len(flying) < cap:
                        item = take(w, not flying)
                        if item is None:
                            break
                        flying.append((item, None))
                        flying[-1] = (item, fn.submit(item[1], item[2]))
                    if not flying:
                        break                    # take() blocked until the stop flag: nothing left to do
                    item, tok = flying[0]
                    res = fn.wait(tok)
                    flying.pop(0)
                except BaseException as exc:  # noqa: BLE001
                    fail(w, exc, [it_ for it_, _ in flying])
                    return
                with cv:
                    done[item[0]] = res
                    cv.notify_all()

        threads = [threading.Thread(target=worker, args=(w, d), daemon=True) for w, d in enumerate(self.devices)]
        for t in threads:
            t.start()
        submitted = emitted = 0
        it = iter(pairs)
        exhausted = False
        try:
            while True:
                while not exhausted and submitted - emitted < self.max_inflight * G:
                    want = self.chunk
                    if self.ramp and submitted < 2 * G:
                        want = max(1, self.chunk // (4 if submitted < G else 2))
                    Ls, Rs = [], []
                    for l, r in it:
                        Ls.append(l); Rs.append(r)
                        if len(Ls) == want:
                            break
                    if not Ls:
                        exhausted = True
                        break
                    with cv:
                        w = submitted % G
                        if dead[w]:              # its turn goes to whoever is free first
                            orphans.append((submitted, Ls, Rs))
                        else:
                            mine[w].append((submitted, Ls, Rs))
                        cv.notify_all()
                    submitted += 1
                    if len(Ls) < want:
                        exhausted = True
                if emitted == submitted and exhausted:
                    break
                with cv:
                    while emitted not in done:
                        if all(dead):
                            raise RuntimeError(f"all GPU workers failed: {errors[-1]!r}") from errors[-1]
                        cv.wait(timeout=0.5)
                    disp, depth = done.pop(emitted)
                emitted += 1
                n = (disp if disp is not None else depth).shape[0]
                for k in range(n):
                    yield (disp[k] if disp is not None else None), (depth[k] if depth is not None else None)
        finally:
            with cv:
                stop[0] = True               # (a consumer that walks away early: what is still queued is dropped)
                orphans.clear()
                for q in mine:
                    q.clear()
                cv.notify_all()
            for t in threads:        # the workers drain what they have in flight: nobody may reuse their engines before that
                t.join(timeout=60.0)
