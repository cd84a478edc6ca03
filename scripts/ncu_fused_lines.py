"""Attribute an ncu report of a specialised (NVRTC-generated) kernel to its generated source lines.
The cubin is regenerated here (NVRTC needs no GPU; generation is deterministic), disassembled with
line info and joined with the report's per-SASS-instruction counters by instruction order.

usage: ncu_fused_lines.py report.ncu-rep workload B T [top]     workload: w3 | w2 | w3b
(B = 0: the pipelined kernel, T = voices per CTA)"""
import collections, csv, glob, io, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
rep, wl, B, T = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
tmp = tempfile.mkdtemp()
os.environ["DUSP_FUSED_DUMP"] = tmp
if B == 0:
    os.environ["DUSP_PIPED"] = "1"
    B = 4
from dusp_b200 import capi, workloads as W
from dusp_b200.render import circuit_blob
import circuits

circ = {"w3": W.w3_circuit, "w2": W.w2_live_circuit, "w3b": W.w3b_circuit}[wl]()
prog = capi.Program(circuit_blob(circ)[0])
prog.specialise(B, T)
cubin = glob.glob(os.path.join(tmp, "*.cubin"))[0]
src = open(glob.glob(os.path.join(tmp, "*.cu"))[0]).read().split("\n")
ops = open(os.path.join(ROOT, "dusp_b200", "csrc", "ops.cuh")).read().split("\n")
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
cur, ins = None, []
for l in sass:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((m.group(2), cur))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
cols, data = rows[hdr], rows[hdr + 1:]
ci, cs = cols.index("Instructions Executed"), cols.index("# Samples")
if len(data) != len(ins):
    print("WARNING: instruction count differs (report %d, regenerated cubin %d): mapping by order is unreliable" % (len(data), len(ins)))
by, bys, byop, byops = collections.Counter(), collections.Counter(), collections.Counter(), collections.Counter()
total = 0
for (txt, c), r in zip(ins, data):
    n, s = int(r[ci]), int(r[cs])
    total += n
    by[c] += n
    bys[c] += s
    op = txt.split()[0] if not txt.startswith("@") else txt.split()[1]
    byop[op] += n
    byops[op] += s
ts = sum(bys.values()) or 1
print("total warp-instructions executed:", total, " static SASS instructions:", len(ins))
for k, n in sorted(by.items(), key=lambda kv: -bys[kv[0]])[:top]:
    f, l = k if k else ("?", 0)
    text = (ops if f == "ops.cuh" else src)
    code = text[l - 1].strip()[:110] if 0 < l <= len(text) else f
    print("%5.1f%% instr %5.1f%% stall-samples  %s:%d  %s" % (100 * n / total, 100 * bys[k] / ts, f, l, code))
print("--- by opcode (instr%, stall-sample%)")
for k, n in byop.most_common(30):
    print("%5.1f%% %5.1f%% %s" % (100 * n / total, 100 * byops[k] / ts, k))
