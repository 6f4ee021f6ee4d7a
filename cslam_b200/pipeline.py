"""Pipelined host-to-host batched search shared by ``NearestNeighborsMatching.search_batches`` and
``ShardedNearestNeighborsMatching.search_batches`` (additive API, SURVEY §8 f1: message batches).

Plumbing only (streams, events, pinned result blocks): the search itself is the stream-ordered device-in /
device-out C-ABI call the owner passes in.  Three streams: uploads (host -> device copy engine), the caller's
current stream (kernels), result reads (device -> host copy engine); up to ``depth`` blocks in flight.
"""
import collections

import numpy as np
import torch


def pipelined_search(owner, batches, k, depth, reads, search_device, items_of, check):
    """Generator of ``(items, sims)`` per block of ``batches`` (host (B, dim) float32 tensors or arrays), in order.
    ``owner`` provides ``dim`` and ``n`` and keeps the streams and staging slots between calls;
    ``search_device(dq, kk, out_idx, out_sim)`` queues one search on the current stream; ``items_of(rows)`` maps
    a (B, kk) row array to item lists; ``check()`` raises on a failed exchange; ``reads`` False = this rank reads
    nothing back and yields ``(None, None)``."""
    k = int(k)
    dev_index = getattr(owner, "_dev_index", None)
    dev = torch.device("cuda", torch.cuda.current_device() if dev_index is None else dev_index)
    compute = torch.cuda.current_stream(dev)
    if getattr(owner, "_pipe_streams", None) is None:
        owner._pipe_streams = (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
        owner._pipe_slots = []
    up, down = owner._pipe_streams
    free = collections.deque(owner._pipe_slots)
    owner._pipe_slots = []
    flight = collections.deque()

    def slot_for(B, dim, kk):
        while free:
            sl = free.popleft()
            if sl["dq"].shape == (B, dim) and sl["idx"].shape == (B, kk):
                return sl
        # a block fresh from torch's caching allocator may still be read by kernels queued earlier on the compute
        # stream (the allocator only orders reuse within one stream): the other two streams catch up first
        up.wait_stream(compute)
        down.wait_stream(compute)
        return {"dq": torch.empty((B, dim), dtype=torch.float32, device=dev),
                "idx": torch.empty((B, kk), dtype=torch.int64, device=dev),
                "sim": torch.empty((B, kk), dtype=torch.float32, device=dev),
                "hidx": torch.empty((B, kk), dtype=torch.int64).pin_memory(),
                "hsim": torch.empty((B, kk), dtype=torch.float32).pin_memory(),
                "up": torch.cuda.Event(), "searched": torch.cuda.Event(), "done": torch.cuda.Event()}

    def submit(block):
        if not torch.is_tensor(block):
            block = torch.from_numpy(np.asarray(block).astype(np.float32))
        q = block.reshape(-1, owner.dim).to(torch.float32)
        B, kk = int(q.shape[0]), min(k, owner.n)
        if kk == 0 or B == 0:
            flight.append((None, B))
            return
        sl = slot_for(B, int(q.shape[1]), kk)
        with torch.cuda.stream(up):          # the slot's last search and read-back are over: see `finish`
            sl["dq"].copy_(q, non_blocking=True)
            sl["up"].record(up)
        compute.wait_event(sl["up"])
        search_device(sl["dq"], kk, sl["idx"], sl["sim"])
        sl["searched"].record(compute)
        if reads:
            with torch.cuda.stream(down):
                down.wait_event(sl["searched"])
                sl["hidx"].copy_(sl["idx"], non_blocking=True)
                sl["hsim"].copy_(sl["sim"], non_blocking=True)
                sl["done"].record(down)
        flight.append((sl, B))

    def finish():
        sl, B = flight.popleft()
        if sl is None:
            return [[] for _ in range(B)], np.empty((B, 0), np.float32)
        (sl["done"] if reads else sl["searched"]).synchronize()
        check()
        out = (None, None)
        if reads:
            out = (items_of(sl["hidx"].numpy()), sl["hsim"].numpy().copy())
        free.append(sl)
        return out

    try:
        for block in batches:
            if len(flight) >= max(1, int(depth)):
                yield finish()
            submit(block)
        while flight:
            yield finish()
    finally:
        while flight:   # the consumer stopped early: let the queued work drain before the slots are reused
            sl, _ = flight.popleft()
            if sl is not None:
                (sl["done"] if reads else sl["searched"]).synchronize()
                free.append(sl)
        owner._pipe_slots = list(free)
