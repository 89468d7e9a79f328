"""Static view of the hot loops of k1_march<false>: python tools/sass_loops.py [lib.so] — lists every backward branch
(loop) with its instruction count and pipe mix (ALU = half-rate pipe on sm_100)."""
import os, re, subprocess, sys, collections
KERN = os.environ.get("KERN", "k1_marchILb0ELi2")
lib = sys.argv[1] if len(sys.argv) > 1 else "gamer_b200/lib/libgamer_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn = None; ins = []
for l in out.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m: fn = m.group(1); continue
    if fn and KERN in fn:
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
ALU = {'LOP3','SHF','IADD3','VIADD','LEA','FSEL','FSETP','FMNMX','SEL','ISETP','MOV','PRMT','PLOP3','VIMNMX','FSET','IABS','FCHK','P2R','R2P','DSETP','FMNMX3','I2FP','F2FP','BMSK','SGXT','POPC','FLO','CS2R','IADD'}
FMA = {'FFMA','FMUL','FADD','IMAD','HFMA2','FFMA2','FMUL2','FADD2'}
def opc(t):
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", t); return m.group(2) if m else "?"
addr2i = {a: i for i, (a, _) in enumerate(ins)}
print("k1_march<false>:", len(ins), "instructions")
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA\S*\s+(?:\S+,\s*)?(0x[0-9a-f]+)", t)
    if m and opc(t) == "BRA":
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in addr2i:
            body = ins[addr2i[tgt]:i + 1]
            c = collections.Counter(opc(x) for _, x in body)
            alu = sum(v for k, v in c.items() if k in ALU); fma = sum(v for k, v in c.items() if k in FMA)
            print(f"loop {tgt:#x}..{a:#x}: {len(body)} instr, ALU {alu}, FMA-pipe {fma}, LDS {c['LDS']}, FP64 {sum(v for k,v in c.items() if k[0]=='D' and k!='DSETP')}, XU {c['MUFU']+c['F2F']+c['F2I']+c['I2F']}, LDL/STL {c['LDL']+c['STL']}  ", dict(c.most_common(8)))
