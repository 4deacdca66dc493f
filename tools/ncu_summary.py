"""Summarise gpurun_out ncu artefacts into profiles/ (tracked).  Usage:
    python tools/ncu_summary.py <tag> <launches.csv> [<full.ncu-rep>]"""
import collections, csv, subprocess, sys, io, json, os
tag, launches = sys.argv[1], sys.argv[2]
rep = sys.argv[3] if len(sys.argv) > 3 else None
out = []
rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
h = rows[0]; ki, vi, ui = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Unit')
agg = collections.OrderedDict()
for r in rows[1:]:
    n = r[ki].split('(')[0]; v = float(r[vi].replace(',', ''))
    v = v / 1e3 if r[ui] == 'ns' else (v * 1e3 if r[ui] == 'ms' else v)
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
out.append("# ncu launch list (%s): --metrics gpu__time_duration.sum --clock-control none; cold-cache serialised, compare SHARES" % tag)
out.append("%-62s %5s %12s %7s" % ("kernel", "n", "us/launch", "share"))
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append("%-62s %5d %12.1f %6.1f%%" % (n[:62], a[0], a[1] / a[0], 100 * a[1] / tot))
traffic = {}
if rep:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(txt)))
    hh, units = rr[0], rr[1]
    want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
            'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
            'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
            'sm__inst_executed_pipe_tensor.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
            'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
            'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct']
    out.append("\n# ncu --set full --clock-control none (%s), one row per captured launch" % tag)
    for r in rr[2:]:
        name = r[hh.index('Kernel Name')].split('(')[0]
        out.append("## " + name)
        vals = {}
        for w in want:
            if w in hh:
                i = hh.index(w); vals[w] = r[i]
                out.append("   %-70s %s %s" % (w, r[i], units[i]))
        try:
            def mb(k):
                i = hh.index(k); v = float(r[i].replace(',', '')); u = units[i]
                return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1)
            traffic[name] = mb('dram__bytes_read.sum') + mb('dram__bytes_write.sum')
        except Exception as e:
            pass
os.makedirs("profiles", exist_ok=True)
open("profiles/%s_ncu_summary.txt" % tag, "w").write("\n".join(out) + "\n")
if traffic:
    json.dump(traffic, open("profiles/%s_traffic_by_kernel.json" % tag, "w"), indent=1)
print("\n".join(out[:12]))
