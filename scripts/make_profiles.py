"""Turns gpurun_out/launches.csv + gpurun_out/prof.ncu-rep (from scripts/prof_final.sh)
into the committed summaries under profiles/."""
import csv, json, re, subprocess, sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
lines = [l for l in open('gpurun_out/launches.csv') if not l.startswith('==')]
rows = list(csv.DictReader(lines))
fmt = lambda g: "x".join(re.findall(r"\d+", g))
names = [(r['Kernel Name'].replace('<unnamed>::', '').split('(')[0].replace('void ', ''), fmt(r['Grid Size']),
          fmt(r['Block Size']), float(r['Metric Value'].replace(',', ''))) for r in rows]
idx_rate = [i for i, x in enumerate(names) if x[0].startswith('k_rate<')]
idx_store = [i for i, x in enumerate(names) if x[0] == 'k_vt_store']


def seq_before(end):
    j = end
    while j > 0 and 'elementwise' not in names[j][0]:
        j -= 1
    return names[j + 1:end + 1]


rounds, stream = seq_before(idx_rate[-1]), seq_before(idx_store[-1])
with open('profiles/%s_launches.csv' % tag, 'w') as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none ; python bench.py --steps 1 --warmup 1 --no-cpu-baseline --skip-e2e\n")
    f.write("# config 2: 10M 3v3 matches over 1M players; the timed step of each path; per-launch times are cold-cache and serialised: compare shares\n")
    for title, seq in (("stream path (headline): ts_link_pred + ts_rate_stream", stream), ("rounds path: ts_schedule + ts_rate", rounds)):
        tot = sum(x[3] for x in seq)
        f.write("# %s -- total %.3f ms\n" % (title, tot / 1e6))
        f.write("kernel,grid,block,duration_us,share\n")
        for n, g, b, v in seq:
            f.write('"%s",%s,%s,%.1f,%.3f\n' % (n, g, b, v / 1e3, v / tot))
out = subprocess.run(['ncu', '-i', 'gpurun_out/prof.ncu-rep', '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum',
        'lts__t_sector_hit_rate.pct']


def tobytes(v, u):
    try:
        v = float(v.replace(',', ''))
    except ValueError:
        return float('nan')
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1)


traffic, seen = {}, set()
with open('profiles/%s_ncu_summary.txt' % tag, 'w') as f:
    f.write("ncu --set full --clock-control none --import-source on, kernels of one bench step (config 2: 10M 3v3 matches, 1M players), B200\n")
    f.write("source: gpurun_out/prof.ncu-rep of scripts/prof_final.sh (not committed: 10+ MB); numbers per launch; first launch of each kernel\n\n")
    for r in rows[2:]:
        name = r[hdr.index('Kernel Name')].replace('<unnamed>::', '').split('(')[0].replace('void ', '')
        if r[hdr.index('dram__bytes_read.sum')].strip() in ('', '-nan', 'nan'):
            continue
        key = name + r[hdr.index('launch__grid_size')]
        if key in seen and not name.startswith('k_radix_onesweep'):
            continue
        seen.add(key)
        f.write("== %s\n" % name)
        for w in want:
            if w in hdr:
                f.write("   %-70s %s %s\n" % (w, r[hdr.index(w)], units[hdr.index(w)]))
        rd = tobytes(r[hdr.index('dram__bytes_read.sum')], units[hdr.index('dram__bytes_read.sum')])
        wr = tobytes(r[hdr.index('dram__bytes_write.sum')], units[hdr.index('dram__bytes_write.sum')])
        k = 'k_rate_stream' if 'k_rate_stream' in name else ('k_rate' if name.startswith('k_rate<') else None)
        if k and k not in traffic:
            traffic[k] = {"matches": 10000000, "slots": 6, "dram_bytes_per_launch": rd + wr, "dram_read": rd,
                          "dram_write": wr, "source": "profiles/%s_ncu_summary.txt" % tag}
json.dump(traffic, open('profiles/rate_kernel_traffic.json', 'w'), indent=1)
print(open('profiles/%s_launches.csv' % tag).read())
print(traffic)
