"""Summarise an ncu report (read here, no GPU needed) into profiles/<name>.txt and update
profiles/traffic.json with the per-launch DRAM traffic of the profiled kernel.

  python tools/ncu_summary.py gpurun_out/gemm_bf16_r1.ncu-rep --name gemm_bf16_r1 \
      [--workload matmul_bf16]
"""

import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KEEP = (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.per_cycle_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__cluster_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
    "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sector_hit_rate.pct",
    "sm__cycles_elapsed.avg", "sm__cycles_active.avg", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "sm__cycles_elapsed.avg.per_second",
)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--name", required=True)
    ap.add_argument("--workload", default="")
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = [f"# ncu --set full --clock-control none, report {os.path.basename(args.report)}",
           "# (metrics taken under the profiler: cold cache, serialised; never a bench value)"]
    traffic = None
    for launch in rows[2:]:
        rec = dict(zip(hdr, launch))
        out.append("")
        out.append(f"kernel: {rec.get('Kernel Name')}  grid {rec.get('Grid Size')} "
                   f"block {rec.get('Block Size')}")
        for h, u, v in zip(hdr, units, launch):
            if h in KEEP:
                out.append(f"  {h:72s} {v} {u}")
        try:
            scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}
            rd = float(rec["dram__bytes_read.sum"]) * scale[units[hdr.index("dram__bytes_read.sum")]]
            wr = float(rec["dram__bytes_write.sum"]) * scale[units[hdr.index("dram__bytes_write.sum")]]
            traffic = rd + wr
            out.append(f"  dram traffic per launch (read+write): {traffic:.0f} bytes")
        except (KeyError, ValueError):
            pass
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", args.name + ".txt"), "w") as f:
        f.write("\n".join(out) + "\n")
    if args.workload and traffic is not None:
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        data = {}
        if os.path.exists(tpath):
            with open(tpath) as f:
                data = json.load(f)
        data[args.workload] = traffic
        with open(tpath, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)
    print("\n".join(out))


if __name__ == "__main__":
    sys.exit(main())
