"""python tools/sass_dump.py lo hi [lib.so] — SASS of k1_march<false> between two addresses (hex)."""
import os, re, subprocess, sys
KERN = os.environ.get("KERN", "k1_marchILb0ELi2")
lo, hi = int(sys.argv[1], 16), int(sys.argv[2], 16)
lib = sys.argv[3] if len(sys.argv) > 3 else "gamer_b200/lib/libgamer_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn = None
for l in out.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m: fn = m.group(1); continue
    if fn and KERN in fn:
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m and lo <= int(m.group(1), 16) <= hi: print(m.group(1), m.group(2).strip())
