"""dusp_mix_push_static (include/dusp_b200.h): the mixdown push for callers that cannot allocate device memory.
No counterpart in the reference.  Collected last on purpose.
Checks the new entry point, through the C ABI, against the existing dusp_mix_push on two calls
in a row, and that sizes beyond the stated maximum are refused."""
import ctypes

import numpy as np
import pytest

from dusp_b200 import capi

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _read(ptr, n, dtype=np.float32):
    import torch
    from dusp_b200.mixdown import _DevArray
    t = torch.as_tensor(_DevArray(ptr, n), device="cuda").cpu().numpy()
    return np.frombuffer(t.tobytes(), dtype=dtype)


def test_static_scratch_push_matches_the_allocating_entry_point(lib):
    import torch
    torch.cuda.set_device(0)
    n = 2 * 5000 + 6  # not a multiple of the kernel's grid stride
    rng = np.random.default_rng(11)
    parts = [rng.uniform(-1, 1, n).astype(np.float32) for _ in range(2)]
    parts[1][::7] = 0.0  # the kernel skips zero addends
    ref = capi.MixAccumulator(0, n, 1)   # outputs are the caller's, for both entry points
    new = capi.MixAccumulator(0, n, 1)
    ticket = torch.zeros(1, dtype=torch.int32, device="cuda")  # the scratch the existing entry point takes from its caller
    st = torch.cuda.current_stream().cuda_stream
    ref.clear(st)
    new.clear(st)
    try:
        for k, part in enumerate(parts):  # two calls in a row: the static ticket keeps the last call's contents
            t = torch.from_numpy(part).cuda()
            ref.push(t.data_ptr(), 0, n, 0, ticket.data_ptr(), st)
            capi.check(lib.dusp_mix_push_static(0, ctypes.c_void_p(t.data_ptr()), ctypes.c_void_p(new.acc), n,
                                                ctypes.c_void_p(new.counters), ctypes.c_void_p(st)))
            torch.cuda.synchronize()
            want, got = _read(ref.acc, n), _read(new.acc, n)
            scale = max(1.0, float(np.max(np.abs(want))))
            assert float(np.max(np.abs(got.astype(np.float64) - want))) <= TOL * scale, "call %d" % k
            assert int(_read(new.counters, 1, np.uint32)[0]) == k + 1 == int(_read(ref.counters, 1, np.uint32)[0])
        assert float(np.max(np.abs(_read(new.acc, n)))) > 0.5  # something was added at all
        # beyond the stated maximum: refused with a message, nothing launched
        rc = lib.dusp_mix_push_static(0, ctypes.c_void_p(ticket.data_ptr()), ctypes.c_void_p(new.acc), (1 << 24) + 1,
                                      ctypes.c_void_p(new.counters), ctypes.c_void_p(st))
        assert rc != 0 and b"maximum" in lib.dusp_last_error()
        torch.cuda.synchronize()
        assert int(_read(new.counters, 1, np.uint32)[0]) == 2
    finally:
        ref.close()
        new.close()
