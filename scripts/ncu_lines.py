"""Attribute an ncu report's per-SASS-instruction counters to CUDA source lines.
usage: ncu_lines.py report.ncu-rep libdusp_b200.so 'ILi16ELi4E' [top]"""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, so, tag = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
cubins = [f for f in os.listdir(tmp) if f.endswith(".cubin")]
cubin = [f for f in cubins if tag in subprocess.run(["cuobjdump", "-elf", os.path.join(tmp, f)], capture_output=True, text=True).stdout][0] if len(cubins) > 1 else cubins[0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.split("\n")
starts = [i for i, l in enumerate(sass) if l.startswith(".text.")]
s0 = [i for i in starts if tag in sass[i]][0]
s1 = min([i for i in starts if i > s0] + [len(sass)])
cur, ins = None, []
for l in sass[s0:s1]:
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
assert len(data) == len(ins), (len(data), len(ins))
by, bys, byop = collections.Counter(), collections.Counter(), collections.Counter()
total = 0
for (txt, c), r in zip(ins, data):
    n, s = int(r[ci]), int(r[cs])
    total += n
    by[c] += n
    bys[c] += s
    byop[txt.split()[0] if not txt.startswith("@") else txt.split()[1]] += n
src = open(os.path.join(os.path.dirname(os.path.abspath(so)), "csrc", "render_sm100.cu")).read().split("\n")
ts = sum(bys.values()) or 1
print("total warp-instructions executed:", total)
for k, n in by.most_common(top):
    f, l = k if k else ("?", 0)
    code = src[l - 1].strip()[:100] if f == "render_sm100.cu" and 0 < l <= len(src) else f
    print("%5.1f%% instr %5.1f%% stall-samples  L%4d %s" % (100 * n / total, 100 * bys[k] / ts, l, code))
print("--- by opcode")
for k, n in byop.most_common(25):
    print("%5.1f%% %s" % (100 * n / total, k))
