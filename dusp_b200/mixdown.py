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

The reduce runs through ``torch.distributed`` (NCCL over NVLink on the GPU
box, gloo in the CPU tests): 2*n floats per time block, latency-bound, issued
asynchronously so that it overlaps the next block's render.
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


class MixdownRenderer:
    """Renders a session block by block in ``DUSP_OUT_MIX`` mode on the current
    CUDA stream and reduces every block's [2][n] mix onto rank 0 while the next
    block renders (two device buffers in rotation).

    ``on_block(t0, n, tensor)`` is called on rank 0 once a block's reduced mix is
    complete (the tensor is recycled afterwards)."""

    def __init__(self, session, block_samples, device):
        import torch
        self.torch = torch
        self.session = session
        self.block = int(block_samples)
        self.bufs = [torch.zeros(2 * self.block, dtype=torch.float32, device=device) for _ in range(2)]
        self.pending = [None, None]  # (work, t0, n) of the reduce in flight on each buffer

    def _finish(self, i, on_block):
        if self.pending[i] is None:
            return
        work, t0, n = self.pending[i]
        if work is not None:
            work.wait()  # orders the current stream after the collective
        if on_block is not None:
            on_block(t0, n, self.bufs[i][:2 * n].view(2, n))
        self.pending[i] = None

    def render(self, n_samples, on_block=None):
        from . import capi
        torch = self.torch
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
