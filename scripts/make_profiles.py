"""Turns the reports of scripts/prof_r2.sh (gpurun_out/r2_launches.csv, r2_prof.ncu-rep, r2_rate.ncu-rep,
r2_big{1,2,3}.ncu-rep, r2_link.ncu-rep) into the committed summaries under profiles/:
  r2_launches.csv            the timed step of both paths, kernel by kernel (shares)
  r2_ncu_summary.txt         ncu --set full numbers of every kernel of a config-2 step
  r2_big_summary.txt         the rating kernels on ladders whose rows do not fit L2, and the link kernels
  r2_big_traffic.json        DRAM bytes per launch at those sizes (bench.py reads it for configs.*.roofline.traffic)
  rate_kernel_traffic.json   the same for the headline kernel at config 2 (bench.py: roofline.traffic)
  r2_sass_k_rate_stream.txt  cuobjdump -sass excerpt of k_rate_stream<6,2,1>: poll loads, row stores, DFMA chain
The .ncu-rep files themselves (8-10 MB each) stay in gpurun_out/."""
import csv, json, os, re, subprocess, sys

tag = "r2"
G = "gpurun_out/"
short = lambda n: n.replace('<unnamed>::', '').replace('void ', '').split('(')[0]

# ---- launch list ----
lines = [l for l in open(G + 'r2_launches.csv') if l.startswith('"')]
rows = list(csv.DictReader(lines))
fmt = lambda g: "x".join(re.findall(r"\d+", g))
names = [(short(r['Kernel Name']), fmt(r['Grid Size']), fmt(r['Block Size']), float(r['Metric Value'].replace(',', ''))) for r in rows]
idx_rate = [i for i, x in enumerate(names) if x[0].startswith('k_rate<')]
idx_store = [i for i, x in enumerate(names) if x[0] == 'k_vt_store']


def seq_before(end):
    j = end
    while j > 0 and 'elementwise' not in names[j][0]:
        j -= 1
    return names[j + 1:end + 1]


rounds, stream = seq_before(idx_rate[-1]), seq_before(idx_store[-1])
with open('profiles/%s_launches.csv' % tag, 'w') as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none ; python bench.py --steps 1 --warmup 1 --no-cpu-baseline --skip-e2e --no-configs\n")
    f.write("# config 2: 10M 3v3 matches over 1M players; the timed step of each path; per-launch times are cold-cache and serialised: compare shares\n")
    for title, seq in (("stream path (headline): ts_link_pred + ts_rate_stream", stream), ("rounds path: ts_schedule + ts_rate", rounds)):
        tot = sum(x[3] for x in seq)
        f.write("# %s -- total %.3f ms\n" % (title, tot / 1e6))
        f.write("kernel,grid,block,duration_us,share\n")
        for n, g, b, v in seq:
            f.write('"%s",%s,%s,%.1f,%.3f\n' % (n, g, b, v / 1e3, v / tot))

want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio', 'smsp__average_warp_latency_issue_stalled_barrier.ratio']


def raw(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return rows[0], rows[1], rows[2:]


def tobytes(v, u):
    try:
        v = float(v.replace(',', ''))
    except ValueError:
        return float('nan')
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1)


def dump(f, hdr, units, r, label=None):
    name = short(r[hdr.index('Kernel Name')])
    f.write("== %s%s\n" % (name, "   [" + label + "]" if label else ""))
    for w in want:
        if w in hdr:
            f.write("   %-72s %s %s\n" % (w, r[hdr.index(w)], units[hdr.index(w)]))
    rd = tobytes(r[hdr.index('dram__bytes_read.sum')], units[hdr.index('dram__bytes_read.sum')])
    wr = tobytes(r[hdr.index('dram__bytes_write.sum')], units[hdr.index('dram__bytes_write.sum')])
    return name, rd, wr


traffic = {}
with open('profiles/%s_ncu_summary.txt' % tag, 'w') as f:
    f.write("ncu --set full --clock-control none --import-source on, kernels of one bench step (config 2: 10M 3v3 matches, 1M players), B200\n")
    f.write("source: gpurun_out/r2_prof.ncu-rep and r2_rate.ncu-rep of scripts/prof_r2.sh (not committed: 10+ MB); numbers per launch\n\n")
    for path in (G + 'r2_rate.ncu-rep', G + 'r2_prof.ncu-rep'):
        if not os.path.exists(path):
            continue
        hdr, units, body = raw(path)
        seen = {}
        for r in body:
            name = short(r[hdr.index('Kernel Name')])
            key = name + r[hdr.index('launch__grid_size')]
            seen[key] = seen.get(key, 0) + 1
            if seen[key] > (3 if name.startswith('k_radix_onesweep') else 1):
                continue
            name, rd, wr = dump(f, hdr, units, r)
            k = 'k_rate_stream' if name.startswith('k_rate_stream') else ('k_rate' if name.startswith('k_rate<') else None)
            if k and k not in traffic:
                traffic[k] = {"matches": 10000000, "slots": 6, "dram_bytes_per_launch": rd + wr, "dram_read": rd,
                              "dram_write": wr, "source": "profiles/%s_ncu_summary.txt" % tag}
json.dump(traffic, open('profiles/rate_kernel_traffic.json', 'w'), indent=1)

big = []
cases = [("r2_big1.ncu-rep", "125M 3v3 matches / 12.5M players, K=1 (config 4's ladder)", 125000000, 6, 1),
         ("r2_big2.ncu-rep", "125M 3v3 matches / 12.5M players, K=2 (shared + queue pair)", 125000000, 6, 2),
         ("r2_big3.ncu-rep", "100M 5v5 matches / 10M players, Zipf, 2% draws (config 3)", 100000000, 10, 1)]
with open('profiles/%s_big_summary.txt' % tag, 'w') as f:
    f.write("ncu --set full --clock-control none --import-source on: streaming rating kernels on ladders whose rows do not fit the 126 MB L2\n")
    f.write("(scripts/prof_big.sh, one launch each), then the link kernels of a 50M-match / 5M-player ladder (scripts/prof_link.py)\n\n")
    for path, label, n, p, k in cases:
        if not os.path.exists(G + path):
            continue
        hdr, units, body = raw(G + path)
        name, rd, wr = dump(f, hdr, units, body[0], label)
        alg = ((4 * p + 4) + 2 * p * 16 * k) * n
        f.write("   DRAM traffic %.1f GB = %.2f x the algorithmic %.1f GB (%d B/match)\n\n" % ((rd + wr) / 1e9, (rd + wr) / alg, alg / 1e9, alg // n))
        big.append({"kernel": "k_rate_stream" if k == 1 else "k_rate_stream_pair", "matches": n, "slots": p, "pairs": k,
                    "dram_bytes_per_launch": rd + wr, "dram_read": rd, "dram_write": wr, "source": "profiles/%s_big_summary.txt" % tag})
    if os.path.exists(G + 'r2_link.ncu-rep'):
        hdr, units, body = raw(G + 'r2_link.ncu-rep')
        seen = {}
        for r in body:
            name = short(r[hdr.index('Kernel Name')])
            seen[name] = seen.get(name, 0) + 1
            if seen[name] <= (2 if name.startswith('k_radix_onesweep') else 1):
                dump(f, hdr, units, r, "link kernels, 50M matches / 5M players (300M slot entries)")
json.dump(big, open('profiles/%s_big_traffic.json' % tag, 'w'), indent=1)

# ---- SASS excerpt of the headline kernel ----
so = "analyzer_b200/libtsrate.so"
if os.path.exists(so):
    sass = subprocess.run(['cuobjdump', '-sass', '-fun', '_ZN41_GLOBAL__N__', so], capture_output=True, text=True).stdout
    if not sass.strip():
        sass = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True).stdout
    # the function of k_rate_stream<6, 2, true>
    m = re.search(r"Function : (\S*k_rate_streamILi6ELi2ELb1E\S*)(.*?)(?=\n\s*Function :|\Z)", sass, re.S)
    if m:
        body = m.group(2).splitlines()
        ops = [re.sub(r"/\*[0-9a-f]+\*/", "", l).strip() for l in body if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        ops = [re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", o) for o in ops if o]
        n_dfma = sum(1 for o in ops if re.search(r"\bD(FMA|MUL|ADD)\b", o))
        keep = [o for o in ops if re.search(r"LDG\.E\.\S*256|STG\.E\.\S*256|LDGSTS|MUFU|NANOSLEEP|VOTE|SHFL|CCTL", o)]
        with open('profiles/%s_sass_k_rate_stream.txt' % tag, 'w') as f:
            f.write("cuobjdump -sass analyzer_b200/libtsrate.so, %s\n" % m.group(1))
            f.write("%d SASS instructions, %d of them DFMA/DMUL/DADD (no HMMA / UTC*MMA: nothing here is a contraction)\n" % (len(ops), n_dfma))
            f.write("memory and synchronisation instructions, in program order (LDG/STG .256 = one access per 32-byte row; STRONG.GPU = relaxed gpu scope):\n")
            for o in keep:
                f.write("    " + o + "\n")
print(open('profiles/%s_launches.csv' % tag).read()[:3000])
print(json.dumps(traffic)); print(json.dumps(big))
