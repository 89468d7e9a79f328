#!/bin/bash
# K2 durations under ncu (each launch on a cold, clean L2; gpu__time_duration is not quantised like the event timing of a
# 15 us kernel): base library and every gamer_b200/lib/variants/lib_k2_*.so, at 2048^2 and 4096^2
mkdir -p gpurun_out
for f in "" gamer_b200/lib/variants/lib_k2_*.so; do
  if [ -z "$f" ]; then unset GAMER_B200_LIB; name=base; else [ -f "$f" ] || continue; export GAMER_B200_LIB=$PWD/$f; name=$(basename $f .so); fi
  for size in 2048 4096; do
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k2_tonemap --csv \
        --log-file gpurun_out/k2ncu_${name}_$size.csv python tools/gpu/k2_loop.py $size 6 > gpurun_out/k2ncu_${name}_$size.log 2>&1
    python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/k2ncu_${name}_$size.csv")) if len(r)>5 and r[0].isdigit()]
d=sorted(float(r[-1]) for r in rows if "time_duration" in r[-3])
rd=[float(r[-1]) for r in rows if "bytes_read" in r[-3]]
print("$name", $size, "us:", " ".join("%.2f"%(x/1000 if x>1000 else x) for x in d), "| dram read", rd[:1], rows[0][-2] if rows else "")
PY
  done
done
