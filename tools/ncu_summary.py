"""Summarise ncu outputs (run in the build container): launch-list CSV -> per-kernel shares; .ncu-rep -> key metrics."""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg",
        "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def launches(path):
    lines = [ln for ln in open(path) if not ln.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
        k = row["Kernel Name"].split("(")[0]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# {sum(v[0] for v in agg.values())} launches, {tot / 1e3:.1f} ms (cold-cache, serialised: compare SHARES)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
        print(f"{k[:72]:72s} n={v[0]:5d} ms={v[1] / 1e3:9.2f} share={v[1] / tot:.3f} avg_us={v[1] / v[0]:8.1f}")


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    hdr = r[0]
    print("# " + path)
    for row in r[2:]:
        d = {h: (row[i], r[1][i]) for i, h in enumerate(hdr)}
        print("kernel:", d["Kernel Name"][0][:100])
        for k in KEYS:
            for h in hdr:
                if h.endswith(k):
                    print(f"   {k:75s} {d[h][0]:>16s} {d[h][1]}")
                    break


if __name__ == "__main__":
    for p in sys.argv[1:]:
        (launches if p.endswith(".csv") else report)(p)
