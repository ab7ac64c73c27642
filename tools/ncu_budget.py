"""Per-stage instruction budget of pair_frames_kernel from an `ncu --page source --csv --print-source cuda,sass` dump.
usage: ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > src.csv ; python tools/ncu_budget.py src.csv <frames per launch> [kernel substring]
Stages are line ranges of chacra_b200/csrc/*.cuh found through anchor strings in the CURRENT sources (run it on the build the
report was taken from); inlined helpers are attributed through their own files."""
import csv
import os
import re
import sys
from collections import defaultdict

CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "chacra_b200", "csrc")
# (stage, file, start anchor, end anchor): from the first line containing `start` up to the line before the first later line containing `end`
ANCHORS = [
    ("prologue / frame setup", "pair_kernel.cuh", "pair_frames_kernel(DevTopo T", "// ---- R1: ring tasks (cation x ring"),
    ("R1 ring prefilter", "pair_kernel.cuh", "// ---- R1: ring tasks (cation x ring", "// the slot-pair arrays: shared memory"),
    ("record -> slot pairs, partner bits", "pair_kernel.cuh", "// the slot-pair arrays: shared memory", "const int px = hdr->nc[0] + 2"),
    ("E state setup", "pair_kernel.cuh", "const int px = hdr->nc[0] + 2", "// Slot pairs, packed float32 arithmetic"),
    ("E1 loop head", "pair_kernel.cuh", "// Slot pairs, packed float32 arithmetic", "k0 += 32;"),
    ("E1 unit grab / ranges", "pair_kernel.cuh", "k0 += 32;", "// up to 16 slot pairs of every lane"),
    ("E1 test loop + filter", "pair_kernel.cuh", "// up to 16 slot pairs of every lane", "// slot k of the chunk sits at bit"),
    ("E1 masks / partner window / entry", "pair_kernel.cuh", "// slot k of the chunk sits at bit", "// ---- E2: 32 filtered survivors"),
    ("E2 dispatch", "pair_kernel.cuh", "// ---- E2: 32 filtered survivors", "// ---- E3: 32 classified pairs"),
    ("E3 dispatch", "pair_kernel.cuh", "// ---- E3: 32 classified pairs", "// ---- R2: the listed ring tasks"),
    ("R2 ring dispatch", "pair_kernel.cuh", "// ---- R2: the listed ring tasks", "// ---- W: water bridges"),
    ("end barrier / water", "pair_kernel.cuh", "// ---- W: water bridges", "\x00"),
    ("queue put/get lambdas", "pair_kernel.cuh", "auto qput = []", "for (int k = tid; k < T.K * T.K"),
    ("slot accessors (E2/E3)", "frame_phases.cuh", "struct FrameCtxT", "// ---- R: pi-cation"),
    ("R2 ring tasks", "frame_phases.cuh", "// ---- R: pi-cation", "// ---- E2: exact float32 d2"),
    ("E2 classify", "frame_phases.cuh", "// ---- E2: exact float32 d2", "// does donor d see acceptor a"),
    ("E3 hb_direction", "frame_phases.cuh", "// does donor d see acceptor a", "// ---- E3: everything that needs"),
    ("E3 finish", "frame_phases.cuh", "// ---- E3: everything that needs", "\x00"),
    ("dist2_exact (E2/E3/R)", "device_types.cuh", "__device__ __forceinline__ float dist2_exact", "// superset prefilter only"),
    ("count_pair", "device_types.cuh", "__device__ __forceinline__ long long pair_index", "\x00"),
    ("mbarrier / bulk copy", "frame_kernel.cuh", "// ---- mbarrier + bulk async copy", "// grid coordinate of x along one axis"),
    ("cell_of", "frame_kernel.cuh", "__device__ __forceinline__ int cell_of", "// ---- output helpers"),
    ("emit records", "frame_kernel.cuh", "// ---- output helpers", "// Per-frame residue-pair set."),
    ("count_pair", "frame_kernel.cuh", "// Per-frame residue-pair set.", "__device__ __forceinline__ bool is_bonded"),
    ("E3 is_bonded / disulfide", "frame_kernel.cuh", "__device__ __forceinline__ bool is_bonded", "// measure hbonds angle"),
    ("R ring float64 math", "frame_kernel.cuh", "// ---- float64 ring geometry", "\x00"),
]


def resolve():
    spec = []
    cache = {}
    for name, fname, a, b in ANCHORS:
        if fname not in cache:
            cache[fname] = open(os.path.join(CSRC, fname)).read()
        text = cache[fname]
        i = text.find(a)
        if i < 0:
            continue
        j = text.find(b, i + len(a)) if b != "\x00" else len(text)
        if j < 0:
            j = len(text)
        lo = text.count("\n", 0, i) + 1
        hi = text.count("\n", 0, j)
        spec.append((name, fname, lo, max(lo, hi)))
    return spec


def main():
    path, frames = sys.argv[1], float(sys.argv[2])
    kern = sys.argv[3] if len(sys.argv) > 3 else "pair_frames_kernel"
    spec = resolve()
    cur = fn = header = None
    acc = defaultdict(lambda: [0.0, 0.0, 0.0])
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "File Path":
            cur = row[1].split("/")[-1]
            continue
        if row[0] == "Function Name":
            fn = row[1]
            continue
        if row[0] == "Line No":
            header = {k: i for i, k in enumerate(row)}
            continue
        if header is None or fn is None or kern not in fn or row[2] not in ("", "-"):
            continue
        try:
            line = int(row[0])
        except ValueError:
            continue
        name = "other:" + cur
        for n, f, lo, hi in spec:
            if f == cur and lo <= line <= hi:
                name = n
                break

        def g(k):
            try:
                return float(row[header[k]] or 0)
            except ValueError:
                return 0.0
        acc[name][0] += g("# Samples"); acc[name][1] += g("Instructions Executed"); acc[name][2] += g("Thread Instructions Executed")
    ts = sum(v[0] for v in acc.values()) or 1
    ti = sum(v[1] for v in acc.values()) or 1
    print("kernel %s: %d samples, %.4g warp instructions (%.0f per frame)" % (kern, ts, ti, ti / frames))
    print("%-34s %8s %8s %11s %8s" % ("stage", "samples%", "winst%", "winst/frame", "thr/inst"))
    for n, v in sorted(acc.items(), key=lambda kv: -kv[1][1]):
        if v[1]:
            print("%-34s %7.1f%% %7.1f%% %11.0f %8.1f" % (n, 100 * v[0] / ts, 100 * v[1] / ti, v[1] / frames, v[2] / v[1]))


if __name__ == "__main__":
    main()
