#!/bin/bash
OUT=gpurun_out
mkdir -p $OUT
for bs in 32 256; do
  python tools/step_timeline.py $bs $OUT/timeline_b$bs.txt > $OUT/a2_timeline_$bs.log 2>&1; echo "timeline b$bs rc=$?"; tail -16 $OUT/timeline_b$bs.txt
  DPB_CONV_CLUSTER_MAX=1 python tools/step_timeline.py $bs $OUT/timeline_b${bs}_nocluster.txt > $OUT/a2_timeline_${bs}_nc.log 2>&1; tail -14 $OUT/timeline_b${bs}_nocluster.txt
done
python -c "from distributed_pytorch_b200 import _native as N; N.build(force=True)" 
echo "== cli tests"
timeout 900 python -m pytest tests/test_gpu_cli.py tests/test_gpu_diagnostics.py -q -p no:cacheprovider 2>&1 | tail -15
echo "== ref"
timeout 300 python bench.py --steps 20 --warmup 5 --impl reference > $OUT/a2_ref.json 2>$OUT/a2_ref.err; cut -c1-300 $OUT/a2_ref.json
echo "== launch lists"
for bs in 32 256; do
  ncu --metrics gpu__time_duration.sum --clock-control none -s 130 -c 70 --csv --log-file $OUT/launches_b$bs.csv python tools/step_profile.py $bs 4 > $OUT/a_ncu_$bs.log 2>&1
  echo "ncu b$bs rc=$?"
done
