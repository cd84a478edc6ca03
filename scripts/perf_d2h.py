"""Host-side ceiling of the end-to-end path: device -> pinned host copy bandwidth per rank, with
1..N ranks copying at the same time (run under torchrun; every rank prints one JSON line, rank 0 a summary).
Variants: pinned by cudaHostAlloc (torch pin_memory) or cudaHostRegister; 1-D copies or the 2-D copies
dusp_session_render issues (rows of one time block into rows of the whole render); 256 MB and 1 GB blocks.
usage: python -m torch.distributed.run --nproc-per-node N scripts/perf_d2h.py   (or plain python for N = 1)"""
import ctypes, json, os, time
import numpy as np
import torch
import torch.distributed as dist

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaMemcpy2DAsync.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t,
                                 ctypes.c_int, ctypes.c_void_p]
rt.cudaHostRegister.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint]
rt.cudaHostUnregister.argtypes = [ctypes.c_void_p]
D2H = 2


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, nbytes, reps=3):
    fn()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    barrier()
    return nbytes * reps / dt / 1e9


res = {"rank": rank, "world": world}
total = 4 << 30  # bytes of host buffer per rank
dev = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for how in ("hostalloc", "hostregister"):
    if how == "hostalloc":
        host = torch.empty(total, dtype=torch.uint8).pin_memory()
        hptr = host.data_ptr()
    else:
        arr = np.empty(total, dtype=np.uint8)
        arr[::4096] = 1  # touch every page first
        hptr = arr.ctypes.data
        rc = rt.cudaHostRegister(ctypes.c_void_p(hptr), total, 0)
        if rc != 0:
            res[how] = "cudaHostRegister failed: %d" % rc
            continue
    for blk in (256 << 20, 1 << 30):
        def one_d():
            for off in range(0, total, blk):
                rt.cudaMemcpy2DAsync(ctypes.c_void_p(hptr + off), blk, ctypes.c_void_p(dev.data_ptr()), blk, blk, 1, D2H, ctypes.c_void_p(st))
        res["%s_1d_%dMB" % (how, blk >> 20)] = timed(one_d, total)
        # 2-D: 20 000 rows (10 000 voices x 2 channels); a block holds blk / 20000 bytes of every row
        rows = 20000
        w = blk // rows // 256 * 256
        pitch = total // rows // 256 * 256
        def two_d():
            for k in range(pitch // w):
                rt.cudaMemcpy2DAsync(ctypes.c_void_p(hptr + k * w), pitch, ctypes.c_void_p(dev.data_ptr()), w, w, rows, D2H, ctypes.c_void_p(st))
        res["%s_2d_%dMB" % (how, blk >> 20)] = timed(two_d, (pitch // w) * w * rows)
    if how == "hostregister":
        rt.cudaHostUnregister(ctypes.c_void_p(hptr))
    else:
        del host
print(json.dumps(res), flush=True)
if world > 1:
    keys = [k for k, v in res.items() if isinstance(v, float)]
    t = torch.tensor([res[k] for k in keys], device="cuda")
    lo = t.clone(); dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    if rank == 0:
        print(json.dumps({"summary": "GB/s per rank (slowest rank) and aggregate over %d ranks" % world,
                          **{k: {"min_per_rank": float(a), "aggregate": float(b)} for k, a, b in zip(keys, lo.tolist(), sm.tolist())}}), flush=True)
    dist.destroy_process_group()
