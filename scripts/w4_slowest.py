"""Which W4 circuits are the tail of the multi-program launch?  (DUSP_MULTI_DEBUG_FILE diagnostics, csrc/capi.cc)
usage: w4_slowest.py rank [n_circuits=2500]      -> ticks per circuit of one 88 064-sample launch, slowest first
       w4_slowest.py only i j ...                -> render just these circuits (two or more), e.g. under ncu"""
import collections, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dusp_b200 import capi, workloads as W
from dusp_b200.render import circuit_blob

mode = sys.argv[1]
blk = 88064
if mode == "rank":
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 2500
    idx = list(range(n))
    path = tempfile.mktemp()
    os.environ["DUSP_MULTI_DEBUG_FILE"] = path
else:
    idx = [int(a) for a in sys.argv[2:]]
progs = [capi.Program(circuit_blob(W.w4_circuit(i))[0]) for i in idx]
s = capi.Session(progs, [1] * len(idx), 0)
buf = torch.empty(len(idx) * blk, dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream()
for rep in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s.render_device(blk, buf.data_ptr(), capi.OUT_VOICES, st.cuda_stream)
    e1.record()
    torch.cuda.synchronize()
    print("launch %d: %.2f ms" % (rep, e0.elapsed_time(e1)))
if mode == "rank":
    rows = [tuple(int(x) for x in l.split()) for l in open(path)]
    last = max(r[0] for r in rows)
    rows = sorted((r for r in rows if r[0] == last), key=lambda r: -r[4])
    tot = sum(r[4] for r in rows)
    print("circuits %d, ticks: max %.1fM mean %.1fM; sum/max = %.1f" % (len(rows), rows[0][4] / 1e6, tot / len(rows) / 1e6, tot / rows[0][4]))
    for r in rows[:12]:
        i = idx[r[2]]
        ops = collections.Counter(l.split()[1] for l in progs[r[2]].disassemble().splitlines()[1:] if l.strip())
        print("circuit %5d  warp %5d  %3d instr  %7.1fM ticks  %s" % (i, r[1], r[3], r[4] / 1e6, dict(ops.most_common())))
    os.unlink(path)
s.close()
