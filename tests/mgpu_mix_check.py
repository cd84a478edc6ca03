"""Multi-GPU check of the mixdown sum (run under torchrun, one process per GPU; see
test_gpu_parity.py::test_mixdown_two_gpus_p2p_matches_nccl_and_single_process): 40 W5 voices sharded by global index,
rendered block-wise with the one-sided NVLink transport and with ncclReduce; rank 0 compares both with one session
holding all voices."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dusp_b200 import capi, mixdown as MD, workloads as W  # noqa: E402
from dusp_b200.render import circuit_blob  # noqa: E402


def session_for(indices, device):
    i2 = [i for i in indices if i % 2 == 0]
    i3 = [i for i in indices if i % 2 == 1]
    b2, _ = circuit_blob(W.w2_live_circuit())
    b3, _ = circuit_blob(W.w3_circuit())
    p2, p3 = capi.Program(b2), capi.Program(b3)
    s = capi.Session([p2, p3], [len(i2), len(i3)], device)
    o2, o3 = W.w2_overrides(b2, W.w2_params(0, 0, i2)), W.w3_overrides(b3, W.w3_params(0, 0, i3))
    s.set_params(0, list(o2), np.stack([o2[k] for k in o2], axis=1))
    s.set_params(1, list(o3), np.stack([o3[k] for k in o3], axis=1))
    return s, (p2, p3)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    total, n, blk = 40, 7000, 2048
    voice0, nv = MD.shard(total, rank, world)
    sess, keep = session_for(range(voice0, voice0 + nv), local)
    results = {}
    for transport in ("p2p", "nccl", "p2p"):  # p2p twice: the accumulator is cleared and reused
        got = np.zeros((2, n), dtype=np.float32)

        def on_block(t0, k, t):
            got[:, t0:t0 + k] = t.cpu().numpy()

        sess.reset()
        r = MD.MixdownRenderer(sess, blk, torch.device("cuda", local), collective=transport)
        r.render(n, on_block)
        if transport == "p2p":
            sess.reset()
            r.render(n, on_block)  # same renderer again
        torch.cuda.synchronize()
        dist.barrier()
        assert r.collective == transport, "transport fell back to %s" % r.collective
        r.close()
        results.setdefault(transport, []).append(got)
    sess.close()
    if rank == 0:
        ref_s, keep2 = session_for(range(total), local)
        whole = ref_s.render_host(n, capi.OUT_MIX)
        ref_s.close()
        scale = max(1.0, float(np.max(np.abs(whole))))
        assert float(np.max(np.abs(whole))) > 0.1
        for transport, gots in results.items():
            for got in gots:
                err = float(np.max(np.abs(got - whole)))
                assert err <= 1e-5 * scale, "%s differs from the single-process mix by %g" % (transport, err)
        print("MIX-CHECK OK", {k: len(v) for k, v in results.items()}, flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
