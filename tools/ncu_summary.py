"""Here (no GPU): condense `ncu -i X.ncu-rep --page raw --csv` into a markdown table of the metrics the roofline uses.
usage: python tools/ncu_summary.py rep1.ncu-rep [rep2 ...] > profiles/rNN_ncu_full_summary.md"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %peak"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor(hmma) %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor inst"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "alu pipe %"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma pipe %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu (MUFU) %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("smsp__cycles_active.avg", "smsp cycles active"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
]


def main():
    print("# ncu --set full summaries (one row per profiled launch; `--clock-control none`)\n")
    for rep in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        ix = {h: i for i, h in enumerate(hdr)}
        print("## %s\n" % rep.split("/")[-1])
        cols = [(k, n) for k, n in KEYS if k in ix]
        print("| # | kernel | " + " | ".join("%s [%s]" % (n, units[ix[k]]) for k, n in cols) + " | dram total / launch |")
        print("|---|---|" + "---|" * (len(cols) + 1))
        for j, r in enumerate(data):
            name = r[ix["Kernel Name"]]
            name = name.replace("b2v::(anonymous namespace)::", "").split("(")[0][:60]
            vals = [r[ix[k]] for k, _ in cols]
            tot = ""
            try:
                rd = float(r[ix["dram__bytes_read.sum"]].replace(",", ""))
                wr = float(r[ix["dram__bytes_write.sum"]].replace(",", ""))
                tot = "%.4g %s" % (rd + wr, units[ix["dram__bytes_read.sum"]]) if units[ix["dram__bytes_read.sum"]] == units[ix["dram__bytes_write.sum"]] else "%s + %s" % (r[ix["dram__bytes_read.sum"]], r[ix["dram__bytes_write.sum"]])
            except Exception:
                pass
            print("| %d | `%s` | " % (j, name) + " | ".join(vals) + " | %s |" % tot)
        print()


if __name__ == "__main__":
    main()
