"""Join an ncu SASS source page (csv) with nvdisasm line info of the relocatable kernel cubin:
per source line of ks_sim_kernel.cu -> executed instructions and stall samples.
usage: ncu_lines.py <src.csv> <kernel.sass from `nvdisasm -g -c`> [top]"""
import csv, re, sys, collections
src_csv, sass, top = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 50
# nvdisasm: instruction offsets with current line
line_of = {}
cur = None; infn = False; stack = []
for ln in open(sass):
    if ln.startswith(".text.ks_replay_kernel:"): infn = True; continue
    if infn and ln.startswith("//-----"): break
    if not infn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
    if m:
        f, l = m.group(1), int(m.group(2))
        cur = (f.split('/')[-1], l, (m.group(3) or '').split('/')[-1], int(m.group(4) or 0)); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*?);', ln)
    if m: line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
hdr = None; recs = []; take = False
for r in rows:
    if r and r[0] == "Kernel Name": take = (r[1] == "ks_replay_kernel"); continue
    if r and r[0] == "Address": hdr = r; continue
    if take and hdr and len(r) > 10: recs.append(dict(zip(hdr, r)))
base = int(recs[0]["Address"], 16)
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
ti = ts = 0
for d in recs:
    off = int(d["Address"], 16) - base
    key = line_of.get(off)
    ins, smp = int(d["Instructions Executed"]), int(d["# Samples"])
    ti += ins; ts += smp
    a = agg[key]; a[0] += ins; a[1] += smp
    for c in stall_cols:
        v = int(d.get(c) or 0)
        if v: a[2][c] += v
print(f"kernel: {ti} instr, {ts} samples, {len(recs)} SASS")
srcs = {}
def text(k):
    if not k: return ""
    f, l = k[0], k[1]
    if f not in srcs:
        import glob
        c = glob.glob(f"/root/repo/**/{f}", recursive=True)
        srcs[f] = open(c[0]).read().splitlines() if c else []
    s = srcs[f]
    return s[l-1].strip()[:90] if 0 < l <= len(s) else ""
for k, (ins, smp, st) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    tops = ",".join(f"{c[6:]}:{100*v//max(smp,1)}" for c, v in st.most_common(3))
    where = f"{k[0]}:{k[1]}" + (f"<-{k[3]}" if k and k[3] else "") if k else "?"
    print(f"{where:34s} ins {100*ins/ti:5.1f}% smp {100*smp/ts:5.1f}% [{tops:42s}] {text(k)}")
