"""Turns one ncu launch list of the wavefront kernels (tools/gpu_profile.sh: gpu__time_duration, dram bytes, smsp__inst_executed,
l1tex__m_xbar2l1tex_read_bytes per launch of `tools/profile_frame.py --spp 16`) plus the ray counts that run printed into
profiles/roofline_inputs.json: warp instructions per ray, L2 -> L1 gather bytes per ray and DRAM bytes per ray of the extend and
shadow kernels. bench.py multiplies them with the rays and times it measures live (CUDA events) to report the issue-rate,
L2-gather and HBM roofs of the traversal kernels.

  python tools/roofline_inputs.py <launches.csv> <plain-run log with the stats dict> <workload> [<traversal_stats.txt>]"""
import ast
import collections
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    csv_path, log_path, workload = sys.argv[1:4]
    rows = list(csv.reader(open(csv_path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H = rows[hdr]
    iK, iM, iV, iID = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("ID")
    L = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) == len(H):
            L.setdefault(r[iID], {"k": r[iK].split("(")[0]})[r[iM]] = float(r[iV].replace(",", ""))
    tot = collections.defaultdict(lambda: collections.defaultdict(float))
    for d in L.values():
        k = "k_wf_shade" if "k_wf_shade" in d["k"] else d["k"]
        for m, v in d.items():
            if m != "k":
                tot[k][m] += v
        tot[k]["launches"] += 1
    stats = ast.literal_eval(re.search(r"\{.*\}", open(log_path).read().strip().splitlines()[-1]).group(0))
    rays = {"k_wf_extend": stats["rays_extend_traced"], "k_wf_shadow": stats["rays_shadow_traced"], "k_wf_shade": stats["rays_extend_traced"]}
    out = {}
    for k, n in rays.items():
        t = tot[k]
        out[k] = {
            "warp_inst_per_ray": t["smsp__inst_executed.sum"] / n,
            "threads_per_inst": t["smsp__thread_inst_executed.sum"] / max(t["smsp__inst_executed.sum"], 1.0),
            "gather_bytes_per_ray": t["l1tex__m_xbar2l1tex_read_bytes.sum"] / n,
            "dram_bytes_per_ray_measured": (t["dram__bytes_read.sum"] + t["dram__bytes_write.sum"]) / n,
            "dram_bytes_per_launch": (t["dram__bytes_read.sum"] + t["dram__bytes_write.sum"]) / max(t["launches"], 1.0),
            "ncu_ms_per_pass": t["gpu__time_duration.sum"] / 1e6,
            "rays_per_pass": n,
            "source": "ncu launch list %s (one 16-sample pass of tools/profile_frame.py, every launch of the kernel summed) / rays the same program traced" % os.path.basename(csv_path),
            "gather_source": "l1tex__m_xbar2l1tex_read_bytes.sum of the same launches: bytes the L1s fetched from the L2 (node and triangle gathers that missed L1, queue reads)",
        }
    out["k_wf_shade"]["algorithmic_bytes_per_hit"] = 16 + 64 + 48 + 64   # hit + ray record in; shadow record + continuation ray out (upper bound); TriShade is scene data
    if len(sys.argv) > 4 and os.path.exists(sys.argv[4]):
        m = re.search(r"nodes/ray ([0-9.]+)\s+tris/ray ([0-9.]+)", open(sys.argv[4]).read())
        if m:
            out["traversal_stats"] = {"node_visits_per_ray": float(m.group(1)), "triangle_tests_per_ray": float(m.group(2)),
                                      "source": "counting build (make -C mobileraytracer_b200/csrc stats, tools/traversal_stats.py), 1 spp frame, extend + shadow rays"}
    path = os.path.join(ROOT, "profiles", "roofline_inputs.json")
    allw = json.load(open(path)) if os.path.exists(path) else {}
    allw[workload] = out
    json.dump(allw, open(path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
