"""Summarise an ncu report (read here, no GPU needed) into the markdown kept under profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--launches gpurun_out/launches.csv] > profiles/rNN_x.md
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "kernel duration"),
    ("sm__cycles_elapsed.max", "SM cycles"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe active %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
]


def ncu(args):
    return subprocess.run(["ncu"] + args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout


def raw_page(rep):
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    return hdr, units, data


def sass_page(rep):
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "sass"]))))
    name = rows[0][1] if rows and rows[0] else "?"
    hdr = rows[1]
    ci = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in rows[2:]:
        if r and r[0] == "Kernel Name":
            break
        if len(r) > ci["# Samples"]:
            data.append(r)
    return name, ci, data


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return 0.0


def main():
    rep = sys.argv[1]
    hdr, units, data = raw_page(rep)
    ci = {h: i for i, h in enumerate(hdr)}
    kn = ci.get("Kernel Name")
    print("# ncu summary: `%s`\n" % rep.split("/")[-1])
    print("Captured with `ncu --set full --clock-control none --import-source on` (per-launch, cold-cache, serialised).\n")
    for row in data[:2]:
        print("## %s\n" % (row[kn] if kn is not None else "kernel"))
        print("| metric | value | unit |\n|---|---|---|")
        for key, label in KEYS:
            if key in ci:
                print("| %s (`%s`) | %s | %s |" % (label, key, row[ci[key]], units[ci[key]]))
        stalls = [(h, num(row[i])) for h, i in ci.items() if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
        stalls.sort(key=lambda t: -t[1])
        print("\nWarp stall reasons (warps per issue-active cycle): " +
              ", ".join("%s %.2f" % (h.split("stalled_")[1].split("_per_issue")[0], v) for h, v in stalls[:7]) + "\n")
    name, sci, sdata = sass_page(rep)
    S, I = sci["# Samples"], sci["Instructions Executed"]
    tot = sum(num(r[S]) for r in sdata) or 1.0
    toti = sum(num(r[I]) for r in sdata) or 1.0
    print("## Where the samples fall (SASS, first captured launch)\n")
    print("Segments between synchronisation instructions; `samples` = share of warp-state samples, `inst` = share of executed warp instructions.\n")
    print("| SASS range | samples | inst | ends with |\n|---|---|---|---|")
    acc = acci = 0.0
    start = 0
    for k, r in enumerate(sdata):
        acc += num(r[S]); acci += num(r[I])
        if any(t in r[1] for t in ("BAR.SYNC", "SYNCS.PHASECHK", "UTCBAR")):
            if acc / tot > 0.004:
                print("| %d..%d | %.1f%% | %.1f%% | `%s` |" % (start, k, 100 * acc / tot, 100 * acci / toti, r[1].strip()[:48]))
            acc = acci = 0.0
            start = k + 1
    ops = collections.Counter()
    for r in sdata:
        t = r[1].strip().split()
        if not t:
            continue
        op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
        ops[op.split(".")[0]] += num(r[I])
    print("\nTop opcodes by executed warp instructions: " + ", ".join("%s %.1f%%" % (o, 100 * n / toti) for o, n in ops.most_common(14)))
    tc = [o for o in ops if o.startswith(("UTC", "LDTM", "STTM", "UTMA", "UBLKCP"))]
    print("\nBlackwell-native opcodes present: " + (", ".join(sorted(tc)) or "none"))
    if "--launches" in sys.argv:
        path = sys.argv[sys.argv.index("--launches") + 1]
        lines = [l for l in open(path) if not l.startswith("==")]
        agg = collections.OrderedDict()
        for row in csv.DictReader(lines):
            try:
                v = float(row["Metric Value"].replace(",", ""))
            except Exception:
                continue
            a = agg.setdefault(row["Kernel Name"][:80], [0, 0.0])
            a[0] += 1; a[1] += v
        t = sum(v[1] for v in agg.values())
        print("\n## Launch list (`--metrics gpu__time_duration.sum`): share of GPU time per kernel\n")
        print("| kernel | launches | total us | share |\n|---|---|---|---|")
        for k, (n, d) in sorted(agg.items(), key=lambda x: -x[1][1])[:14]:
            print("| `%s` | %d | %.1f | %.1f%% |" % (k, n, d / 1e3, 100 * d / t))


if __name__ == "__main__":
    main()
