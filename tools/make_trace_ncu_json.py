#!/usr/bin/env python
"""profiles/trace_kernel_ncu.json from `ncu --set full` captures of the per-ray traversal launches of one pass
(tools/run_traffic.sh writes gpurun_out/traffic_<workload>.ncu-rep; needs ncu to READ the reports, no GPU):

    python tools/make_trace_ncu_json.py c3=gpurun_out/traffic_c3.ncu-rep c4=gpurun_out/traffic_c4.ncu-rep

Per workload: DRAM bytes per ray over all captured closest-hit launches (bench.py's `roofline.traffic`), and the counters
of the LONGEST launch (bounce 1: the one that decides the step) -- issue-slot utilisation, lanes per instruction, alu /
fma pipe, L1 / L2 hit rates, DRAM throughput.  The rays of each launch come from the per-bounce ray counts the probe
prints (gpurun_out/traffic_<workload>.log, `rays per trace launch`), else from the bench line given as rays=<n>.
The file is stamped with bench.trace_build_hash() (sources + flags of the traversal kernel) and the git commit."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

M = {"dur_us": "gpu__time_duration.sum", "inst": "smsp__inst_executed.sum",
     "issue_slot_utilisation_pct": "sm__inst_issued.avg.pct_of_peak_sustained_active",
     "alu_pipe_pct": "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
     "fma_pipe_pct": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
     "lanes_per_instruction": "smsp__thread_inst_executed_per_inst_executed.ratio",
     "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
     "l1_hit_pct": "l1tex__t_sector_hit_rate.pct", "l2_hit_pct": "lts__t_sector_hit_rate.pct",
     "dram_read": "dram__bytes_read.sum", "dram_write": "dram__bytes_write.sum",
     "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
     "l2_sectors": "lts__t_sectors.sum", "l1_sectors": "l1tex__t_sectors.sum", "l1_miss_sectors": "l1tex__t_sectors_lookup_miss.sum"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6,
        "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    res = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")]}
        for k, name in M.items():
            i = next((j for j, h in enumerate(hdr) if h == name or h.endswith("." + name)), None)  # some carry a section prefix
            if i is not None and r[i] != "":
                d[k] = float(r[i].replace(",", "")) * UNIT.get(units[i], 1.0)
        res.append(d)
    return res


def spheres(rep):
    """profiles/spheres_kernel_ncu.json from a capture of render_spheres_kernel (tools/run_c1.sh)."""
    r = next(x for x in rows_of(rep) if "render_spheres" in x["kernel"])
    out = {"_source": "tools/run_c1.sh (ncu --set full --clock-control none -k regex:render_spheres -c 1 on `bench.py --workload c1`) "
                      "-> tools/make_trace_ncu_json.py spheres=<rep>: one 16-frame launch of the configs[0] scene",
           "build_hash": bench.kernel_build_hash(bench.SPHERES_SOURCES),
           "commit": subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip(),
           "c1": {"launch_us": r["dur_us"], "warp_instructions": r["inst"], "dram_bytes_per_launch": r["dram_read"] + r["dram_write"],
                  **{k: r[k] for k in ("issue_slot_utilisation_pct", "lanes_per_instruction", "alu_pipe_pct", "fma_pipe_pct",
                                       "warps_active_pct", "l1_hit_pct", "l2_hit_pct", "dram_pct_of_peak") if k in r}}}
    json.dump(out, open(os.path.join(ROOT, "profiles", "spheres_kernel_ncu.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


def main():
    args = []
    for a in sys.argv[1:]:
        if a.startswith("spheres="):
            spheres(a.split("=", 1)[1])
        else:
            args.append(a)
    if not args:
        return
    sys.argv[1:] = args
    out = {"_source": "tools/run_traffic.sh (ncu --set full --clock-control none, -k regex:trace_kernel, one pass of "
                      "tools/packet_probe.py with camera-ray packets on) -> tools/make_trace_ncu_json.py. dram_bytes_per_ray = "
                      "(dram__bytes_read.sum + dram__bytes_write.sum) summed over the captured closest-hit launches / the rays "
                      "they traced; the other figures are those of the longest launch (bounce 1).",
           "build_hash": bench.trace_build_hash(),
           "commit": subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()}
    for arg in sys.argv[1:]:
        name, rep = arg.split("=", 1)
        rows = [r for r in rows_of(rep) if "trace_kernel" in r["kernel"]]
        log = os.path.splitext(rep)[0] + ".log"
        rays = None
        if os.path.exists(log):
            for line in open(log):
                if line.startswith("closest-hit rays traced by the per-ray kernel in one pass:"):
                    rays = float(line.split(":")[1])
        if not rays:
            raise SystemExit(f"{log}: no ray count (tools/packet_probe.py prints it)")
        dram = sum(r["dram_read"] + r["dram_write"] for r in rows)
        top = max(rows, key=lambda r: r["dur_us"])
        out[name] = {"dram_bytes_per_ray": dram / rays, "launches_captured": len(rows), "rays_traced": rays,
                     "longest_launch_us": top["dur_us"], "warp_instructions_longest_launch": top["inst"],
                     **{k: top[k] for k in ("issue_slot_utilisation_pct", "lanes_per_instruction", "alu_pipe_pct", "fma_pipe_pct",
                                            "warps_active_pct", "l1_hit_pct", "l2_hit_pct", "dram_pct_of_peak") if k in top}}
        # achieved bandwidth at each level during the longest launch (32-byte sectors): what the loads ask of L1, what L1
        # passes on to L2 (its misses; reads), all L2 traffic (incl. stores and the other die's fabric), and DRAM
        sec = top["dur_us"] * 1e-6
        if "l1_sectors" in top:
            out[name]["l1_request_GBps"] = top["l1_sectors"] * 32.0 / sec / 1e9
        if "l1_miss_sectors" in top:
            out[name]["l1_to_l2_GBps"] = top["l1_miss_sectors"] * 32.0 / sec / 1e9
        if "l2_sectors" in top:
            out[name]["l2_GBps"] = top["l2_sectors"] * 32.0 / sec / 1e9
        out[name]["dram_GBps"] = (top["dram_read"] + top["dram_write"]) / sec / 1e9
    path = os.path.join(ROOT, "profiles", "trace_kernel_ncu.json")
    if os.path.exists(path):  # keep the workloads of an earlier capture of the SAME build
        old = json.load(open(path))
        if old.get("build_hash") == out["build_hash"]:
            for k, v in old.items():
                if isinstance(v, dict) and k not in out:
                    out[k] = v
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
