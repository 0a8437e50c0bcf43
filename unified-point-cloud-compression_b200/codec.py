"""Frame / block sharding driver: one process per GPU, no data-path collective.

The codec is intra-only and the reference already codes the blocks of a frame independently and serially
(/root/reference/model/model.py:137-176, :225-238, no halo), so a sequence shards by frame and a frame by block.
Units are assigned largest-first (LPT) from their point counts -- a deterministic function of (sizes, world), so
every rank computes the same schedule without talking to the others.  Results (symbol buffers, reconstructions,
timings) are gathered on the host with `torch.distributed.gather_object`; nothing crosses NVLink.
"""
import torch
import torch.distributed as dist


def lpt_schedule(sizes, world):
    """Largest-processing-time-first assignment of units to `world` ranks.  Returns owner[i] for every unit.
    Ties are broken by unit index, loads by rank index: the same answer on every rank."""
    order = sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i))
    load = [0] * world
    owner = [0] * len(sizes)
    for i in order:
        r = min(range(world), key=lambda j: (load[j], j))
        owner[i] = r
        load[r] += int(sizes[i])
    return owner


def my_units(sizes, world, rank):
    owner = lpt_schedule(sizes, world)
    return [i for i, r in enumerate(owner) if r == rank]


def gather_results(local, world, rank, dst=0):
    """local: {unit index: picklable result}.  Returns the merged dict on `dst`, None elsewhere."""
    if world == 1:
        return dict(local)
    bucket = [None] * world if rank == dst else None
    dist.gather_object(local, bucket, dst=dst)
    if rank != dst:
        return None
    merged = {}
    for part in bucket:
        merged.update(part)
    return merged


class StreamCodec:
    """Codes the frames of a sequence that fall to this rank.  `net` is a UnifiedModel on this rank's GPU.

    coder="host": compress + decompress in the component form (model.py:94-250), strings through the host range coder.
    coder="bypass": the same GPU work per block (`encode_block` / `decode_block`) with the symbols handed from the encoder to the
    decoder on the device -- the metric's definition (the range coder stays on the host and outside the graft).
    Blocks follow the reference's partition (model.py:121-127; evaluate.py:39-46: block_size 1024 for 10-bit, 512 for 11-bit)."""

    def __init__(self, net, q, block_size=1024, rank=0, world=1, coder="host", batched=False):
        """batched (bypass coder): the blocks of a frame are coded as ONE batched sparse tensor (block = batch item) instead of
        one after the other -- same per-block results, one set of launches per frame instead of one per block."""
        assert coder in ("host", "bypass")
        self.net, self.q, self.block_size = net, q, block_size
        self.rank, self.world, self.coder, self.batched = rank, world, coder, batched

    def code_frame(self, cloud):
        """compress + decompress one frame -> (N,6) reconstruction."""
        with torch.no_grad():
            if self.coder == "host":
                strings, shapes, ks, coords, qs = self.net.compress(cloud, self.q, block_size=self.block_size)
                return self.net.decompress(coordinates=coords, strings=strings, shape=shapes, k=ks, q_vals=qs)
            from .model.model import partition_blocks
            x_sorted, counts = partition_blocks(cloud, self.block_size)
            if self.batched and len(counts) > 1:
                enc = self.net.encode_blocks(x_sorted, counts, self.q)
                ys = enc["y_symbols"]
                x_hat = self.net.decode_block(enc["y_points"], enc["z_symbols"], lambda idx: ys, enc["k"], enc["q"])
                return self.net.finalize(x_hat.C, x_hat.F)
            pts, feats, start = [], [], 0
            for count in counts:
                enc = self.net.encode_block(x_sorted[start:start + count], self.q)
                ys = enc["y_symbols"]
                x_hat = self.net.decode_block(enc["y_points"], enc["z_symbols"], lambda idx: ys, enc["k"], self.q)
                pts.append(x_hat.C); feats.append(x_hat.F)
                start += count
            return self.net.finalize(torch.cat(pts, 0), torch.cat(feats, 0))

    def run(self, frames, sizes=None, keep=lambda rec: int(rec.shape[0]), pipeline=None):
        """frames: sequence of (N,6) tensors or zero-argument loaders; sizes: point counts used for the schedule
        (default: len of each frame).  pipeline: a FramePipeline -> this rank's frames are coded `depth` at a time.
        Returns {frame index: keep(reconstruction)} gathered on rank 0."""
        if sizes is None:
            sizes = [int(f.shape[0]) for f in frames]
        mine = my_units(sizes, self.world, self.rank)
        load = lambda i: frames[i]() if callable(frames[i]) else frames[i]
        if pipeline is None:
            local = {i: keep(self.code_frame(load(i))) for i in mine}
        else:
            recs = pipeline.map(lambda i: self.code_frame(load(i)), mine)
            local = {i: keep(r) for i, r in zip(mine, recs)}
        return gather_results(local, self.world, self.rank)


def record_stream(obj, stream):
    """`tensor.record_stream(stream)` on every CUDA tensor reachable from a worker's result (tensors, lists / tuples / dicts of
    them, SparseTensor-like objects with `.F` / `.C`): the consumer's stream now also owns the blocks."""
    if torch.is_tensor(obj):
        if obj.is_cuda:
            obj.record_stream(stream)
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            record_stream(o, stream)
    elif isinstance(obj, dict):
        for o in obj.values():
            record_stream(o, stream)
    elif hasattr(obj, "F") and hasattr(obj, "C"):
        record_stream(obj.F, stream)
        record_stream(obj.C, stream)


class FramePipeline:
    """Several frames in flight on ONE GPU: `depth` host threads, each with its own CUDA stream.

    Coding a frame is a chain of ~450 launches whose buffer sizes are data dependent (coordinate-set sizes come back
    from the device: ~18 stream synchronisations per frame), so a single stream leaves the GPU idle for ~10 % of a
    frame while the host wakes up and enqueues again.  Frames are independent (intra-only codec), so a second frame on
    another stream fills those gaps; the ctypes calls into libpcg.so, torch's launches and `.item()` all release the
    GIL.  Every worker owns its scratch workspace (sparse_ops.scratch is keyed by stream) and its tensors.  The caches the
    workers share -- packed weights, packed prior parameters, host CDF tables -- are filled by `model.prepare()` here, before
    any worker starts, and a later miss (a parameter changed) is taken under a lock with the packing stream synchronised
    before the buffer is published (sparse_ops.pack_lock).  Tensors handed back to the caller are `record_stream`-ed on the
    caller's stream, so the caching allocator does not recycle their blocks on the worker stream while the caller still
    reads them.  No data-path collective, no cross-stream dependency other than the optional `after` event of `submit`.
    """

    def __init__(self, device, depth=2, model=None):
        import queue
        import threading
        self.device = torch.device(device)
        self.depth = int(depth)
        if model is not None:
            model.prepare()
        self._jobs = queue.Queue()
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(self.depth)]
        self._threads = [threading.Thread(target=self._work, args=(s,), daemon=True) for s in self.streams]
        for t in self._threads:
            t.start()

    def _work(self, stream):
        torch.cuda.set_device(self.device)
        with torch.cuda.stream(stream), torch.no_grad():
            while True:
                job = self._jobs.get()
                if job is None:
                    return
                fn, args, after, fut = job
                try:
                    if after is not None:
                        stream.wait_event(after)
                    res = fn(*args)
                    done = torch.cuda.Event()
                    done.record(stream)
                    fut.set_result((res, done))
                except BaseException as e:          # surfaced by Future.result() in the submitting thread
                    fut.set_exception(e)

    def submit(self, fn, *args, after=None):
        """Run fn(*args) on the next free worker (its stream current, no_grad).  `after`: a CUDA event the worker's
        stream waits for first.  Returns a Future of (result, CUDA event recorded when the worker's launches end); a caller
        that consumes the result on another stream waits for the event and calls `record_stream(result, that_stream)`."""
        from concurrent.futures import Future
        fut = Future()
        self._jobs.put((fn, args, after, fut))
        return fut

    def map(self, fn, items, after=None):
        """fn(item) for every item, `depth` at a time; results in order.  The calling stream waits for all of them."""
        futs = [self.submit(fn, it, after=after) for it in items]
        out = []
        cur = torch.cuda.current_stream(self.device)
        for f in futs:
            res, done = f.result()
            cur.wait_event(done)
            record_stream(res, cur)
            out.append(res)
        return out

    def close(self):
        for _ in self._threads:
            self._jobs.put(None)
        for t in self._threads:
            t.join()
