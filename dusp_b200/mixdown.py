"""Multi-GPU sharding of a voice batch and the one collective of the path: the
sum-reduce of per-GPU stereo mixdowns.

Voices are independent circuits (``Inlet.connect`` refuses cross-circuit
wiring, src/Inlet.js:57-58), so a batch shards by contiguous ranges of the
GLOBAL voice index, one process per GPU, with no data-path collective for
per-voice output.  In ``DUSP_OUT_MIX`` mode every GPU accumulates its voices
into ``mix[2][n]`` inside the render kernel; the per-GPU mixes are then summed
onto rank 0 -- what a left-leaning ``Sum.many`` chain over all voices
(src/components/Sum.js:18-29) followed by the copy-out would produce, up to
float summation order (tolerance-based parity, 1e-5 of the mix's full scale).

The sum has two transports (``MixdownRenderer``): one-sided adds into rank 0's
accumulator over NVLink peer memory (``csrc/mixpush.cu``, the default on a GPU
box) or ``torch.distributed.reduce`` (NCCL; gloo in the CPU tests): 2*n floats
per time block either way, overlapping the next block's render.
"""


def shard(n_total, rank, world):
    """Contiguous range of global voice indices owned by ``rank``: (voice0, n).
    Ranges differ by at most one voice and cover [0, n_total) exactly."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    base, extra = divmod(int(n_total), world)
    n = base + (1 if rank < extra else 0)
    voice0 = rank * base + min(rank, extra)
    return voice0, n


def reduce_mix(mix, dst=0, group=None, async_op=False):
    """Sum the per-rank stereo mixes ``mix`` (a [2][n] float32 tensor, CPU or
    CUDA) onto rank ``dst``.  No-op without an initialised process group."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None
    return dist.reduce(mix, dst=dst, op=dist.ReduceOp.SUM, group=group, async_op=async_op)


class _DevArray:
    """A raw device address as a ``__cuda_array_interface__`` object (float32, 1-D) so torch can view it."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False), "version": 3,
                                         "strides": None}


class MixdownRenderer:
    """Renders a session block by block in ``DUSP_OUT_MIX`` mode on the current CUDA stream and sums every
    block's [2][n] mix over all ranks onto rank 0 while the next block renders.

    Two transports for the sum (``collective``):

    * ``"p2p"`` (default with several CUDA ranks on one box): one-sided.  Rank 0 owns an accumulator
      [n_blocks][2][block] that every process maps through CUDA IPC; after rendering a block each rank adds it
      there with system-scope float reductions over NVLink and bumps the block's arrival counter
      (``csrc/mixpush.cu``).  No rank ever waits for a peer; rank 0 consumes block b (one block late) once its
      counter reads ``world``.  One host barrier per ``render`` call orders the accumulator's clearing.
    * ``"nccl"``: ``torch.distributed.reduce`` per block, asynchronous (also the fallback when IPC mapping is
      refused, and what the gloo CPU tests exercise through ``reduce_mix``).

    ``on_block(t0, n, tensor)`` is called on rank 0 once a block's summed mix is complete (the tensor is
    recycled afterwards)."""

    def __init__(self, session, block_samples, device, collective="auto"):
        import torch
        import torch.distributed as dist
        self.torch = torch
        self.session = session
        self.block = int(block_samples)
        self.device = device
        self.bufs = [torch.zeros(2 * self.block, dtype=torch.float32, device=device) for _ in range(2)]
        self.pending = [None, None]  # (work, t0, n) of the reduce in flight on each buffer
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        want_p2p = self.world > 1 and collective in ("auto", "p2p") and torch.device(device).type == "cuda"
        self.collective = "p2p" if want_p2p else ("nccl" if self.world > 1 else "none")
        self.acc = None          # capi.MixAccumulator
        self.acc_blocks = 0
        self.ticket = torch.zeros(1, dtype=torch.int32, device=device) if want_p2p else None

    # ---- one-sided transport
    def _ensure_acc(self, n_blocks):
        """(Re)create rank 0's accumulator for ``n_blocks`` blocks and map it everywhere; falls back to NCCL for
        good if any rank cannot map it."""
        import torch.distributed as dist
        from . import capi
        torch = self.torch
        if self.acc is not None and self.acc_blocks >= n_blocks:
            return True
        if self.acc is not None:
            torch.cuda.synchronize()
            dist.barrier()
            self.acc.close()
            self.acc = None
        dev = torch.device(self.device).index or 0
        n_floats = n_blocks * 2 * self.block
        ok, handle = 1, [None]
        try:
            if self.rank == 0:
                self.acc = capi.MixAccumulator(dev, n_floats, n_blocks)
                handle = [self.acc.handle]
        except capi.DuspError:
            ok, handle = 0, [None]
        dist.broadcast_object_list(handle, src=0)
        if self.rank != 0:
            try:
                if handle[0] is None:
                    raise capi.DuspError("root could not create the accumulator")
                self.acc = capi.MixAccumulator(dev, n_floats, n_blocks, handle=handle[0])
            except capi.DuspError:
                ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            if self.acc is not None:
                self.acc.close()
                self.acc = None
            self.collective = "nccl"
            return False
        self.acc_blocks = n_blocks
        return True

    def _render_p2p(self, n_samples, on_block):
        import torch.distributed as dist
        from . import capi
        torch = self.torch
        stream = torch.cuda.current_stream()
        sp = stream.cuda_stream
        blocks = [(t0, min(self.block, int(n_samples) - t0)) for t0 in range(0, int(n_samples), self.block)]
        if self.rank == 0:
            self.acc.clear(sp)
            stream.synchronize()
        dist.barrier()  # nobody adds into the accumulator before it is cleared

        def consume(b):
            t0, n = blocks[b]
            self.acc.wait(b, self.world, sp)  # holds the stream until every rank's block b has landed
            if on_block is not None:
                view = torch.as_tensor(_DevArray(self.acc.acc + 4 * b * 2 * self.block, 2 * n), device=self.device)
                on_block(t0, n, view.view(2, n))

        for b, (t0, n) in enumerate(blocks):
            buf = self.bufs[b & 1]
            self.session.render_device(n, buf.data_ptr(), capi.OUT_MIX, sp)
            self.acc.push(buf.data_ptr(), b * 2 * self.block, 2 * n, b, self.ticket.data_ptr(), sp)
            if self.rank == 0 and b >= 1:
                consume(b - 1)  # one block late: the peers have had a whole block's time to arrive
        if self.rank == 0:
            consume(len(blocks) - 1)
            flag = torch.as_tensor(_DevArray(self.acc.timed_out_ptr, 1), device=self.device)  # 4 bytes, bit pattern of a u32
            if int(flag.view(torch.int32).item()) != 0:
                raise capi.DuspError("mixdown: a rank's block never arrived at the accumulator (waited 30 s)")

    def close(self):
        if self.acc is not None:
            self.torch.cuda.synchronize()
            self.acc.close()
            self.acc = None

    # ---- NCCL / gloo transport
    def _finish(self, i, on_block):
        if self.pending[i] is None:
            return
        work, t0, n = self.pending[i]
        if work is not None:
            work.wait()  # orders the current stream after the collective
        if on_block is not None and self.rank == 0:  # the other ranks' buffers are undefined after a reduce onto rank 0
            on_block(t0, n, self.bufs[i][:2 * n].view(2, n))
        self.pending[i] = None

    def render(self, n_samples, on_block=None):
        from . import capi
        torch = self.torch
        if self.collective == "p2p":
            n_blocks = (int(n_samples) + self.block - 1) // self.block
            if self._ensure_acc(n_blocks):
                return self._render_p2p(n_samples, on_block)
        stream = torch.cuda.current_stream()
        i = 0
        for t0 in range(0, int(n_samples), self.block):
            n = min(self.block, int(n_samples) - t0)
            self._finish(i, on_block)  # buffer i is free again
            buf = self.bufs[i]
            self.session.render_device(n, buf.data_ptr(), capi.OUT_MIX, stream.cuda_stream)
            work = reduce_mix(buf[:2 * n], dst=0, async_op=True)
            self.pending[i] = (work, t0, n)
            i ^= 1
        self._finish(i, on_block)
        self._finish(i ^ 1, on_block)
