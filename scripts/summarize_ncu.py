"""Summarise gpurun_out ncu artefacts into profiles/ (tracked).  Usage:
   python scripts/summarize_ncu.py <tag> [launches.csv] [name=report.ncu-rep ...]"""
import collections, csv, os, re, subprocess, sys

KEYS = ["gpu__time_duration.sum", "gpc__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "sm__cycles_elapsed.avg",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "lts__t_bytes.sum",
        "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "lts__t_sectors_srcunit_ltcfabric.sum", "lts__t_sector_hit_rate.pct"]


def launch_table(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        name = re.sub(r"\(.*", "", r[ki])[:70]
        agg[name][0] += 1
        agg[name][1] += float(r[vi].replace(",", "")) / 1e6
    tot = sum(v[1] for v in agg.values())
    out = ["| kernel | launches | total ms | share |", "|---|---:|---:|---:|"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
        out.append(f"| `{k}` | {v[0]} | {v[1]:.2f} | {100 * v[1] / tot:.2f} % |")
    out.append(f"| all ({len(rows) - 1} launches) | | {tot:.1f} | 100 % |")
    return "\n".join(out)


def report_table(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, u = rows[0], rows[1]
    out = [f"kernel: `{rows[2][h.index('Kernel Name')][:90]}`  grid {rows[2][h.index('Grid Size')]} block {rows[2][h.index('Block Size')]}", "",
           "| metric | unit | " + " | ".join(f"launch {i}" for i in range(len(rows) - 2)) + " |", "|---|---|" + "---:|" * (len(rows) - 2)]
    for k in KEYS:
        if k in h:
            i = h.index(k)
            out.append(f"| {k} | {u[i]} | " + " | ".join(r[i] for r in rows[2:]) + " |")
    return "\n".join(out)


if __name__ == "__main__":
    tag = sys.argv[1]
    os.makedirs("profiles", exist_ok=True)
    parts = [f"# ncu summary {tag}", ""]
    for a in sys.argv[2:]:
        if a.endswith(".csv"):
            parts += [f"## launch list (`{a}`; gpu__time_duration.sum, --clock-control none; cold-cache, serialised: compare shares)", "", launch_table(a), ""]
        else:
            name, path = a.split("=")
            parts += [f"## {name} (`ncu --set full`, `{path}`)", "", report_table(path), ""]
    open(os.path.join("profiles", tag + ".md"), "w").write("\n".join(parts))
    print("\n".join(parts))
