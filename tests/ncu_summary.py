"""Summarise an `ncu --page raw --csv` export: the handful of metrics the roofline discussion uses."""
import csv
import sys

WANT = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_sector_hit_rate.pct', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio']


def main(path, out=None, note=""):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(w) for w in WANT if w in hdr]
    lines = []
    if note:
        lines.append("# " + note)
    lines.append(",".join("%s [%s]" % (hdr[i], units[i]) for i in idx))
    for r in rows[2:]:
        lines.append(",".join('"%s"' % r[i] for i in idx))
    text = "\n".join(lines) + "\n"
    if out:
        open(out, "w").write(text)
    return rows, hdr, idx


if __name__ == "__main__":
    rows, hdr, idx = main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None,
                          sys.argv[3] if len(sys.argv) > 3 else "")
    for r in rows[2:]:
        print("---", r[hdr.index('Kernel Name')][:60])
        for i in idx[1:]:
            print("   %-85s %s" % (hdr[i], r[i]))
