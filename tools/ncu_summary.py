import csv, collections, subprocess, sys
rep = sys.argv[1]; pts = float(sys.argv[2])
raw = subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines())); hdr=rows[0]
keys=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','sm__cycles_elapsed.max','smsp__average_warp_latency_issue_stalled_long_scoreboard.pct','smsp__warps_eligible.avg.per_cycle_active','smsp__warps_active.avg.per_cycle_active']
for r in rows[2:3]:
    for k in keys:
        if k in hdr: print(k, r[hdr.index(k)], rows[1][hdr.index(k)])
src = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","sass"],capture_output=True,text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = rows[1]
isrc, iex, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [i for i,h in enumerate(hdr) if h.startswith("stall_")]
ops = collections.Counter(); samp = collections.Counter(); tot=0; stalls=collections.Counter()
for r in rows[2:]:
    if len(r) < len(hdr): continue
    if r[0].startswith("Kernel"): break
    toks = r[isrc].split()
    if not toks: continue
    op = toks[0] if not toks[0].startswith('@') else toks[1]
    op = op.split('.')[0]
    n = int(r[iex]); ops[op]+=n; samp[op]+=int(r[isamp]); tot+=n
    for i in stall_cols:
        try: stalls[hdr[i]] += int(r[i])
        except: pass
print("dynamic warp instr", tot, "per warp-point", tot/pts)
for op,n in ops.most_common(28): print(f"{op:10s} {n/pts:8.2f} samples {samp[op]}")
print(stalls.most_common(12))
