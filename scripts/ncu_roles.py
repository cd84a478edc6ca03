"""Per-warp-role stall attribution of an ncu report of the pipelined kernel (source page, SASS view).
Roles are told apart by the generated source line a SASS instruction maps to (nvdisasm -g of the dumped cubin):
P-warps / S-warp / F-warp regions of dusp_fused_*.cu; inlined ops.cuh code inherits the role of the nearest
preceding generated-source line in address order.
usage: ncu_roles.py report.ncu-rep dumped.cubin dumped.cu"""
import collections, csv, io, os, re, subprocess, sys
rep, cubin, cu = sys.argv[1:4]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open(cu).read().split("\n")
ops = open(os.path.join(ROOT, "dusp_b200", "csrc", "ops.cuh")).read().split("\n")
marks = {}
for i, l in enumerate(src, 1):
    if "P-warps: lanes along time" in l: marks["P"] = i
    if "S-warp: one voice per lane" in l: marks["S"] = i
    if "F-warp: one voice per lane" in l: marks["F"] = i
    if "---------------- P-warps" in l and "P" not in marks: marks["P"] = i
order = sorted(marks.items(), key=lambda kv: kv[1])


def role_of(line):
    r = "pre"
    for name, ln in order:
        if line >= ln:
            r = name
    return r


sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
ins, cur_role, cur = [], "pre", None
for l in sass:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        f, ln = m.group(1).split("/")[-1], int(m.group(2))
        cur = (f, ln)
        if f.startswith("dusp_fused"):
            cur_role = role_of(ln)
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((m.group(2), cur, cur_role))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
cols, data = rows[hdr], rows[hdr + 1:]
if len(data) != len(ins):
    print("WARNING: instruction count differs (report %d, cubin %d)" % (len(data), len(ins)))
ci, cs = cols.index("Instructions Executed"), cols.index("# Samples")
stall_cols = [i for i, c in enumerate(cols) if c.startswith("stall_") and "Not Issued" not in c]
per = collections.defaultdict(collections.Counter)
by = collections.defaultdict(collections.Counter)
for (txt, c, role), r in zip(ins, data):
    per[role]["instr"] += int(r[ci] or 0)
    per[role]["samples"] += int(r[cs] or 0)
    for i in stall_cols:
        try:
            per[role][cols[i]] += int(r[i] or 0)
        except ValueError:
            pass
    by[role][(c, txt.split()[0] if not txt.startswith("@") else txt.split()[1])] += int(r[cs] or 0)
for role, cnt in per.items():
    tot = cnt["samples"] or 1
    print("== %s: %d warp-instructions, %d samples" % (role, cnt["instr"], cnt["samples"]))
    for k, v in cnt.most_common(10):
        if k not in ("instr", "samples") and v:
            print("    %-28s %6.1f %%" % (k, 100.0 * v / tot))
for role in by:
    print("-- top SASS sites of", role)
    for (c, op), v in by[role].most_common(10):
        f, ln = c if c else ("?", 0)
        text = ""
        try:
            text = (src if f.startswith("dusp_fused") else ops)[ln - 1].strip()[:90]
        except Exception:
            pass
        print("   %6d  %-12s %s:%d  %s" % (v, op, f, ln, text))
