"""Print the handful of ncu metrics that matter for the sweep kernels from a .ncu-rep (development aid)."""
import csv
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'sm__cycles_elapsed.avg.per_second', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'lts__t_bytes.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__warps_active.avg.per_cycle_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = dict(zip(hdr, vals))
        print("kernel:", d.get("Kernel Name"))
        for h, u in zip(hdr, units):
            if h in WANT:
                print(f"  {h:75s} {d[h]} {u}")
        # warp-state section of --set full: smsp__average_warps_issue_stalled_<reason>_per_issue_active.ratio (warp cycles
        # spent in <reason> per issued instruction); older ncu versions name it ..._per_warp_active.pct
        stalls = {}
        for h in hdr:
            if "issue_stalled_" not in h or not d.get(h):
                continue
            if not (h.endswith("_per_issue_active.ratio") or h.endswith("_per_warp_active.pct")):
                continue
            reason = h.split("issue_stalled_")[1].rsplit("_per_", 1)[0]
            try:
                stalls[(reason, h.rsplit(".", 1)[1])] = float(d[h].replace(",", ""))
            except ValueError:
                pass
        top = sorted(stalls.items(), key=lambda x: -x[1])[:10]
        print("  stalls:", ", ".join(f"{r} {v:.2f} {'cyc/inst' if kind == 'ratio' else '%'}" for (r, kind), v in top))
        if not top:
            print("  (no warp-state metrics matched; headers mentioning 'stall':", [h for h in hdr if "stall" in h][:12], ")")


if __name__ == "__main__":
    main(sys.argv[1])
