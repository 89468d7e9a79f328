"""Summarise an ncu report per source line: python tools/ncu_lines.py rep.ncu-rep samples octaves [top] [files]

Every SASS instruction is counted ONCE, keyed by its address: the source page lists an instruction of inlined code under
each level of its inline stack, and summing those rows over-counts (the round-1 file did: 18.2 G against the 15.1 G of
smsp__inst_executed.sum). An instruction is attributed to the first source line that lists it (the innermost frame).
The total printed here must equal smsp__inst_executed.sum of the same report."""
import collections, csv, re, subprocess, sys
rep, samples, octaves = sys.argv[1], float(sys.argv[2]) / 32, float(sys.argv[3]) / 32
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, vals = rows[0], rows[2]
m = dict(zip(hdr, vals))
for k in ("gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
          "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
          "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
          "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
          "dram__bytes_read.sum", "dram__bytes_write.sum"):
    print(k, m.get(k))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur = None; line = None
text = {}; seen = {}
for r in csv.reader(src.splitlines()):
    if len(r) >= 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) > 7 and r[0].isdigit():
        line = (cur, int(r[0])); text[line] = r[1].strip()[:100]
        continue
    if len(r) > 7 and r[0] == '' and r[2].startswith('0x'):
        addr = int(r[2], 16)
        if addr in seen: continue
        mm = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[3])
        op = mm.group(2).split('.')[0] if mm else '?'
        try: n = int(r[7]); st = int(r[4])
        except ValueError: continue
        seen[addr] = (line, op, n, st, r[3].strip())
lines = collections.Counter(); stall = collections.Counter(); ops = collections.defaultdict(collections.Counter)
for addr, (ln, op, n, st, _) in seen.items():
    lines[ln] += n; stall[ln] += st; ops[ln[0]][op] += n
tot = sum(lines.values()); ts = sum(stall.values())
byfile = collections.Counter()
for (f, l), v in lines.items(): byfile[f] += v
print("total warp-instr (each SASS address once)", tot, {f: round(100 * v / tot, 1) for f, v in byfile.items()})
print("warp-instr per ray sample:", round(tot / samples, 1), " in noise.cuh per octave:", round(byfile['noise.cuh'] / octaves, 1),
      " outside noise.cuh per sample:", round((tot - byfile['noise.cuh']) / samples, 1))
print("noise opcodes per octave:", {k: round(v / octaves, 1) for k, v in ops['noise.cuh'].most_common(16)})
for (f, l), v in lines.most_common(int(sys.argv[4]) if len(sys.argv) > 4 else 45):
    print(f"{f:11s} {l:4d} {100*v/tot:5.1f}% stall {100*stall[(f,l)]/max(ts,1):5.1f}%  {text[(f,l)]}")
if len(sys.argv) > 5:        # extra: every line of the named files in line order, warp-instr per sample
    for f in sys.argv[5].split(','):
        print("----", f, "(warp-instr per ray sample)")
        for (ff, l), v in sorted(lines.items()):
            if ff == f and v / samples >= 0.05:
                print(f"{l:4d} {v/samples:7.2f}  {text[(ff,l)]}")
if len(sys.argv) > 6:        # SASS addresses by execution count: python ... top files addr
    base = min(seen)
    for addr in sorted(seen):
        ln, op, n, st, txt = seen[addr]
        if n / samples >= float(sys.argv[6]): print(f"{addr-base:6x} {n/samples:7.2f} st {100*st/max(ts,1):5.2f}  {ln[0]}:{ln[1]}  {txt[:70]}")
