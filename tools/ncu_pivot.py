"""Pivot an `ncu --csv` metrics log (one row per launch x metric) into one row per launch.
usage: python tools/ncu_pivot.py gpurun_out/engine_launches.csv [--top N]"""
import csv, sys, collections
path = sys.argv[1]
rows = collections.OrderedDict()
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    k = int(r["ID"])
    d = rows.setdefault(k, {"name": r["Kernel Name"].split("(")[0], "grid": r["Grid Size"], "block": r["Block Size"]})
    try:
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    except ValueError:
        d[r["Metric Name"]] = r["Metric Value"]
    d[r["Metric Name"] + "_unit"] = r["Metric Unit"]
def dur_us(d):
    v = d.get("gpu__time_duration.sum", 0.0); u = d.get("gpu__time_duration.sum_unit", "ns")
    return v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
def byt(d, k):
    v = d.get(k, 0.0); u = d.get(k + "_unit", "byte")
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
if __name__ == "__main__":
    tot = sum(dur_us(d) for d in rows.values())
    print("launches %d total %.1f ms" % (len(rows), tot / 1e3))
    by = collections.defaultdict(lambda: [0, 0.0])
    for d in rows.values():
        by[d["name"]][0] += 1; by[d["name"]][1] += dur_us(d)
    for n, (c, t) in sorted(by.items(), key=lambda x: -x[1][1]):
        print("%-48s n=%5d  %9.2f ms  %5.1f%%" % (n[:48], c, t / 1e3, 100 * t / tot))
    if "--detail" in sys.argv:
        for k, d in rows.items():
            print("%4d %-28s grid %-14s %9.1f us  tensor%% %5.1f  dramR %8.1f MB dramW %8.1f MB  L2 %9.1f MB" % (
                k, d["name"][:28], d["grid"], dur_us(d), d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", -1),
                byt(d, "dram__bytes_read.sum") / 1e6, byt(d, "dram__bytes_write.sum") / 1e6, byt(d, "lts__t_bytes.sum") / 1e6))
    if "--dram" in sys.argv:
        print("\nper-kernel DRAM traffic and achieved bandwidth (ncu: dram__bytes_read.sum + dram__bytes_write.sum over "
              "gpu__time_duration.sum; cold-cache, serialised)")
        agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
        for d in rows.values():
            a = agg[d["name"]]
            a[0] += 1; a[1] += dur_us(d); a[2] += byt(d, "dram__bytes_read.sum"); a[3] += byt(d, "dram__bytes_write.sum")
        for n, (c, t, r, w) in sorted(agg.items(), key=lambda x: -x[1][1]):
            if t < 50: continue
            print("%-62s n=%4d %8.2f ms  R %7.2f GB W %7.2f GB  -> %6.0f GB/s  (%.1f MB/launch)" % (
                n[:62], c, t / 1e3, r / 1e9, w / 1e9, (r + w) / (t * 1e-6) / 1e9, (r + w) / c / 1e6))
