"""Where the time goes, by code region: python tools/ncu_regions.py rep.ncu-rep
For each region: share of executed warp instructions, share of the warp-state samples (time), their ratio (relative
cycles per instruction; 1 = the kernel's average) and the mix of stall reasons among the region's samples.
Each SASS address is counted once (see ncu_lines.py)."""
import collections, csv, subprocess, sys
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
hdr = None; seen = {}; cur = None; line = None
for r in csv.reader(src.splitlines()):
    if len(r) >= 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r and r[0] == 'Line No': hdr = r; continue
    if len(r) > 7 and r[0].isdigit(): line = (cur, int(r[0]), r[1]); continue
    if len(r) > 7 and r[0] == '' and r[2].startswith('0x'):
        a = int(r[2], 16)
        if a not in seen: seen[a] = (line, r)
ix = {n: i for i, n in enumerate(hdr)}
import re
MARK = {}      # kernels.cu line -> region, from "// @region name" comments is overkill: use the function the line sits in
def region(ln):
    f, l, text = ln
    if f == 'noise.cuh': return 'noise (octaves, skew)'
    if f in ('dmath.cuh', 'qtf.cuh'): return f + ' (FP64 exp/atan/sqrt/div; twirl)'
    if f == 'march.cuh': return 'march.cuh (MUFU helpers)'
    if f != 'kernels.cu': return 'other'
    return KREG(l)
ksrc = open('gamer_b200/csrc/kernels.cu').read().splitlines()
starts = []
for i, t in enumerate(ksrc, 1):
    for pat, name in (('void arm_base(', 'arm_base'), ('float find_difference(', 'arm_value'), ('void bulge_add(', 'bulge_add'),
                      ('float stars_helper_octaves(', 'stars_helper'), ('void eval_component(', 'eval_component (gates, class code)'),
                      ('void march(', 'march: ray set-up'), ('// ---- run of off-slab samples', 'march: run loop'),
                      ('if (need) {   // ---- one in-slab sample', 'march: in-slab frame'), ('int image_row(', 'kernel: tile loop, ray, clip'),
                      ('// ---- K2:', 'other')):
        if pat in t: starts.append((i, name))
def KREG(l):
    name = 'other'
    for i, n in starts:
        if i <= l: name = n
    return name
S = collections.Counter(); N = collections.Counter(); st = collections.defaultdict(collections.Counter)
keys = ['stall_wait', 'stall_short_sb', 'stall_long_sb', 'stall_math', 'stall_no_inst', 'stall_branch_resolving', 'stall_dispatch',
        'stall_not_selected', 'stall_selected', 'stall_mio']
for a, (ln, r) in seen.items():
    try: n = int(r[7]); s = int(r[6])
    except ValueError: continue
    g = region(ln); S[g] += s; N[g] += n
    for k in keys:
        try: st[g][k] += int(r[ix[k]])
        except ValueError: pass
totS = sum(S.values()); totN = sum(N.values())
print(f"total warp instructions {totN}, samples {totS}")
print(f"{'region':48s} instr%  time%  rel.CPI | stall mix, % of the region's samples")
for g in sorted(S, key=lambda g: -S[g]):
    mix = ' '.join(f"{k[6:]}:{100*st[g][k]/max(S[g],1):.0f}" for k in keys)
    print(f"{g:48s} {100*N[g]/totN:5.1f} {100*S[g]/totS:6.1f}  {(S[g]/totS)/max(N[g]/totN,1e-9):6.2f} | {mix}")
