"""Condense an .ncu-rep (read with `ncu -i … --page raw --csv`) into the handful of metrics the roofline discussion
needs; writes markdown to stdout.   python tools/ncu_summary.py gpurun_out/x.ncu-rep [title]"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__cycles_elapsed.max", "SM cycles"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "memory throughput %"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "DRAM bytes read"),
    ("dram__bytes_write.sum", "DRAM bytes written"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
    ("sm__inst_executed_pipe_tc.sum", "tensor-pipe instructions"),
    ("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (alt)"),
    ("sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed", "HMMA pipe active %"),
    ("sm__inst_executed.sum", "instructions"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__cluster_size", "cluster size"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
    ("smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio", "stall sleeping"),
]


def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[0]
    units = rows[1] if len(rows) > 1 else []
    print(f"## {title}\n")
    for r in rows[2:]:
        rec = dict(zip(hdr, r))
        print(f"**{rec.get('Kernel Name', '?')[:110]}**  (id {rec.get('ID')})\n")
        print("| metric | value | unit |\n|---|---|---|")
        for key, name in KEYS:
            if key in rec and rec[key] != "":
                u = units[hdr.index(key)] if units else ""
                print(f"| {name} (`{key}`) | {rec[key]} | {u} |")
        tc = [k for k in hdr if k.split(".")[0].startswith(("sm__", "l1tex__", "TPC", "smsp__")) and ("_tc_" in k or "tensor" in k) and k not in dict(KEYS) and rec.get(k) not in ("", "0", None)]
        for k in tc[:12]:
            print(f"| `{k}` | {rec[k]} | {units[hdr.index(k)] if units else ''} |")
        print()


if __name__ == "__main__":
    main()
