#!/bin/bash
# ncu --set full captures, converted to CSV on the box (the .ncu-rep files together exceed what gpurun copies back).
OUT=gpurun_out
mkdir -p $OUT
M='dram__bytes_read.sum|dram__bytes_write.sum|dram__throughput|gpu__dram_throughput|gpu__time_duration.sum|sm__pipe_tensor|sm__warps_active|launch__registers_per_thread|launch__grid_size|launch__block_size|lts__t_bytes|l1tex__t_bytes|sm__throughput|lts__throughput|smsp__cycles_active|sm__cycles_elapsed.avg.per_second|launch__cluster'
conv() { ncu -i $OUT/$1.ncu-rep --page raw --csv > $OUT/$1_raw.csv 2>/dev/null; python - <<PY
import csv
rows=list(csv.reader(open("$OUT/$1_raw.csv")))
hdr=rows[0]; import re
keep=[i for i,h in enumerate(hdr) if i<12 or re.search(r"$M",h)]
with open("$OUT/$1.csv","w",newline="") as f:
    w=csv.writer(f)
    for r in rows: w.writerow([r[i] for i in keep if i<len(r)])
print("$1", len(rows)-2, "kernels", len(keep), "columns")
PY
rm -f $OUT/$1_raw.csv; }
python tools/ncu_comm.py 8 > $OUT/ncu_comm_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:"allreduce|gather_mean|broadcast" -s 4 -c 4 -o $OUT/ncu_comm -f python tools/ncu_comm.py 8 > $OUT/ncu_comm.log 2>&1
echo "ncu comm rc=$?"; conv ncu_comm; rm -f $OUT/ncu_comm.ncu-rep
python tools/step_profile.py 256 3 > $OUT/ncu_step_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:"bn_|sgd_kernel|splitk|head_|first_conv|column_reduce" -s 31 -c 31 -o $OUT/ncu_membound -f python tools/step_profile.py 256 3 > $OUT/ncu_membound.log 2>&1
echo "ncu membound rc=$?"; conv ncu_membound; rm -f $OUT/ncu_membound.ncu-rep
python tools/step_profile.py 256 3 > /dev/null 2>&1 && \
ncu --set full --clock-control none -k regex:"conv_gemm" -s 21 -c 21 -o $OUT/ncu_conv -f python tools/step_profile.py 256 3 > $OUT/ncu_conv.log 2>&1
echo "ncu conv rc=$?"; conv ncu_conv; rm -f $OUT/ncu_conv.ncu-rep
for bs in 32 256; do
  ncu --metrics gpu__time_duration.sum --clock-control none -s 110 -c 60 --csv --log-file $OUT/launches_b$bs.csv python tools/step_profile.py $bs 4 > $OUT/a_ncu_$bs.log 2>&1
done
du -sh $OUT
