#!/usr/bin/env python
"""Summarise an `ncu --set full` report as the metric,unit,value CSV kept under profiles/ (one kernel launch).

  python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/ncu_full_rNN_name.csv [--top-stalls 12] [--launch N]
(--launch N: only the N-th launch of the report, counted from 0 -- one file per kernel)
Reads the report with `ncu -i ... --page raw --csv` (and `--page source --csv` for the most-sampled instructions)."""
import csv
import io
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second", "launch__grid_size",
    "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__cluster_size",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed",
    "l1tex__m_xbar2l1tex_read_bytes.sum.per_second", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warp_latency_per_inst_issued.ratio", "smsp__inst_executed.sum", "smsp__inst_executed_op_tma_ld.sum",
]


def run(args):
    return subprocess.run(args, capture_output=True, text=True, check=True).stdout


def main():
    rep, out = sys.argv[1], sys.argv[2]
    top = int(sys.argv[sys.argv.index("--top-stalls") + 1]) if "--top-stalls" in sys.argv else 10
    pick = ["--launch-skip", sys.argv[sys.argv.index("--launch") + 1], "--launch-count", "1"] if "--launch" in sys.argv else []
    rows = list(csv.reader(io.StringIO(run(["ncu", "-i", rep, "--page", "raw", "--csv"] + pick))))
    hdr, units = rows[0], rows[1]
    lines = ["metric,unit,value"]
    for launch in rows[2:]:
        rec = dict(zip(hdr, launch))
        lines.append(f'Kernel Name,,"{rec.get("Kernel Name", "")}"')
        for h, u in zip(hdr, units):
            if h in KEEP and rec.get(h, "") != "":
                lines.append(f"{h},{u},{rec[h]}")
    try:
        src = list(csv.reader(io.StringIO(run(["ncu", "-i", rep, "--page", "source", "--csv"] + pick))))
        h = src[1]
        ia, isrc, isamp, iex = h.index("Address"), h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
        data = [(int(r[isamp] or 0), r[ia][-5:], r[isrc].strip(), r[iex]) for r in src[2:] if len(r) > isamp and (r[isamp] or "0").isdigit()]
        tot = sum(d[0] for d in data) or 1
        lines.append(f"stall_samples_total,sample,{tot}")
        for smp, addr, ins, ex in sorted(set(data), reverse=True)[:top]:
            lines.append(f'"most sampled: {addr} {ins[:70]} (executed {ex})",pct_of_samples,{100.0 * smp / tot:.2f}')
    except Exception as e:  # noqa: BLE001
        lines.append(f'"source page unavailable: {e}",,')
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
