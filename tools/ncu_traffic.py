"""profiles/ncu_traffic.json from an `ncu --set full` capture of tools/profile_run.py: measured DRAM bytes per frame of the two kernels,
tagged with the kernel build (chc_build_tag) so that bench.py never attaches the figure to another build.
usage: ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_traffic.py raw.csv <frames per launch> <build tag> <source name>"""
import csv
import json
import os
import sys


def main():
    path, frames, tag, source = sys.argv[1], float(sys.argv[2]), sys.argv[3], sys.argv[4]
    rows = list(csv.reader(open(path)))
    head = rows[0]
    units = rows[1]
    col = {k: i for i, k in enumerate(head)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    out = {"build_tag": tag, "source": source, "frames_per_launch": frames}
    acc = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        key = "pair_frames_kernel" if "pair_frames_kernel" in name else ("bin_frames_kernel" if "bin_frames" in name else None)
        if key is None:
            continue
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[col[m]].replace(",", "")) * scale[units[col[m]]]
        acc.setdefault(key, []).append(tot)
    for k, v in acc.items():
        out[k + "_bytes_per_frame"] = sum(v) / len(v) / frames
        out[k + "_launches"] = len(v)
    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
