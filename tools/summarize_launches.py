"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of `python bench.py ...` (run HERE, no GPU).

    python tools/summarize_launches.py gpurun_out/r1_launches.csv profiles/r01_launches --warmup 3 --steps 1

Finds the timed region of bench.py from the cfg_ddim_step_kernel launches (one per frame window per video step: the
timed region starts after the last launch of warm-up step W and ends with the last cfg_ddim launch of step W+K) and
writes
  <out>_timed_step.csv   one row per launch in the timed region (id, short kernel name, grid, block, ns)
  <out>_summary.md       per-kernel totals / shares of the timed region and of the whole run
ncu serialises launches and runs them cold-cache, so only the SHARES are meaningful (B200_PROFILING.md).
"""
import argparse
import collections
import csv
import re


def short(name):
    name = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", name)
    name = re.sub(r"^void ", "", name)
    m = re.match(r"([\w:]+(?:<[^()]*>)?)", name)
    return (m.group(1) if m else name)[:90]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("out")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--windows", type=int, default=2, help="frame windows per video step on this rank")
    a = ap.parse_args()
    with open(a.csv) as f:
        lines = [l for l in f if l.startswith('"')]
    rows = []
    for d in csv.DictReader(lines):
        rows.append((int(d["ID"]), short(d["Kernel Name"]), d["Grid Size"], d["Block Size"], float(d["Metric Value"])))
    steps = [i for i, r in enumerate(rows) if "cfg_ddim_step" in r[1]]
    lo = steps[a.warmup * a.windows - 1] + 1
    hi = steps[(a.warmup + a.steps) * a.windows - 1] + 1
    timed = rows[lo:hi]
    with open(a.out + "_timed_step.csv", "w") as f:
        f.write("id,kernel,grid,block,ns\n")
        for r in timed:
            f.write('%d,"%s","%s","%s",%d\n' % (r[0], r[1], r[2], r[3], r[4]))

    def table(rs):
        tot = sum(r[4] for r in rs)
        agg = collections.defaultdict(lambda: [0, 0.0])
        for r in rs:
            agg[r[1]][0] += 1
            agg[r[1]][1] += r[4]
        out = ["| kernel | launches | total ms | share |", "|---|---:|---:|---:|"]
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            out.append("| `%s` | %d | %.3f | %.1f %% |" % (k, n, t / 1e6, 100 * t / tot))
        out.append("| **total** | %d | %.3f | 100 %% |" % (len(rs), tot / 1e6))
        return "\n".join(out)

    with open(a.out + "_summary.md", "w") as f:
        f.write("# ncu launch list summary (%s)\n\n" % a.csv)
        f.write("Command: `ncu --metrics gpu__time_duration.sum --clock-control none --csv python bench.py --steps %d "
                "--warmup %d --no-cpu-baseline`. Launches are serialised and cold-cache under ncu: compare shares, not "
                "absolutes.\n\n" % (a.steps, a.warmup))
        f.write("## Timed region (launch ids %d..%d: %d video step(s), %d frame windows each)\n\n" %
                (timed[0][0], timed[-1][0], a.steps, a.windows))
        f.write(table(timed) + "\n\n## Whole run (weight packing, warm-up, timed step, e2e leg, Merkle leg)\n\n")
        f.write(table(rows) + "\n")
    print(open(a.out + "_summary.md").read())


if __name__ == "__main__":
    main()
