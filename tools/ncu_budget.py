"""Per-stage instruction budget of pair_frames_kernel from an `ncu --page source --csv --print-source cuda,sass` dump.
usage: ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > src.csv ; python tools/ncu_budget.py src.csv pair_frames_kernel <frames per launch>
The line ranges below name the stages of chacra_b200/csrc/pair_kernel.cuh (E1 test loop / flush / unit grab / F / E2 / E3 / ring / prologue)
for the build the report was taken from; inlined helpers are attributed through their own files."""
import csv, sys
from collections import defaultdict
path = sys.argv[1]; kern = sys.argv[2]; frames = float(sys.argv[3])
SHIFT = int(sys.argv[4]) if len(sys.argv) > 4 else 0     # report line = current line + SHIFT in pair_kernel.cuh (r01_v6 capture: -4)
spec = [
 ("prologue+frame setup", "pair_kernel.cuh", 57, 157),
 ("E1 flush (scan+ffs+qput)", "pair_kernel.cuh", 158, 182),
 ("E1 unit grab/ranges", "pair_kernel.cuh", 183, 216),
 ("E1 test loop", "pair_kernel.cuh", 217, 237),
 ("F filter", "pair_kernel.cuh", 238, 257),
 ("E2 dispatch", "pair_kernel.cuh", 258, 279),
 ("E3 dispatch", "pair_kernel.cuh", 280, 295),
 ("R ring dispatch", "pair_kernel.cuh", 296, 321),
 ("end barrier/water", "pair_kernel.cuh", 322, 331),
 ("qput/qget lambdas", "pair_kernel.cuh", 74, 79),
 ("E1 test loop", "device_types.cuh", 105, 109),
 ("dist2_exact (E2/E3/R)", "device_types.cuh", 99, 103),
 ("count_pair idx", "device_types.cuh", 111, 114),
 ("R ring tasks", "frame_phases.cuh", 18, 122),
 ("F filter", "frame_phases.cuh", 124, 133),
 ("E2 classify", "frame_phases.cuh", 135, 177),
 ("E3 hb_direction", "frame_phases.cuh", 179, 206),
 ("E3 finish", "frame_phases.cuh", 208, 278),
 ("cell_of", "frame_kernel.cuh", 56, 59),
 ("emit_contact/bridge", "frame_kernel.cuh", 61, 85),
 ("emit_near", "frame_kernel.cuh", 101, 107),
 ("count_pair", "frame_kernel.cuh", 109, 133),
 ("E3 is_bonded", "frame_kernel.cuh", 135, 139),
 ("E3 is_disulfide", "frame_kernel.cuh", 141, 147),
 ("R ring f64 math", "frame_kernel.cuh", 160, 172),
 ("mbarrier/bulk", "frame_kernel.cuh", 13, 31),
]
cur=None; fn=None; header=None
acc = defaultdict(lambda:[0.0,0.0,0.0])
for row in csv.reader(open(path)):
    if not row: continue
    if row[0]=="File Path": cur=row[1].split("/")[-1]; continue
    if row[0]=="Function Name": fn=row[1]; continue
    if row[0]=="Line No": header={k:i for i,k in enumerate(row)}; continue
    if header is None or kern not in fn: continue
    if row[2] not in ("","-"): continue
    try: line=int(row[0])
    except ValueError: continue
    name="other:"+cur
    for n,f,lo,hi in spec:
        sh = SHIFT if f == "pair_kernel.cuh" else 0
        if f==cur and lo+sh<=line<=hi+sh: name=n; break
    def g(k):
        try: return float(row[header[k]] or 0)
        except ValueError: return 0.0
    acc[name][0]+=g("# Samples"); acc[name][1]+=g("Instructions Executed"); acc[name][2]+=g("Thread Instructions Executed")
ts=sum(v[0] for v in acc.values()) or 1; ti=sum(v[1] for v in acc.values()) or 1
print("kernel %s: %d samples, %.4g warp instructions (%.0f per frame)" % (kern, ts, ti, ti/frames))
print("%-28s %8s %8s %10s %8s" % ("stage","samples%","winst%","winst/frame","thr/inst"))
for n,v in sorted(acc.items(), key=lambda kv:-kv[1][1]):
    print("%-28s %7.1f%% %7.1f%% %10.0f %8.1f" % (n,100*v[0]/ts,100*v[1]/ti,v[1]/frames,v[2]/v[1] if v[1] else 0))
