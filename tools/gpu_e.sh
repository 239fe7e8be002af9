#!/bin/bash
OUT=gpurun_out
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider --junitxml=$OUT/gpu_tests.xml > $OUT/e_pytest.log 2>&1
echo "pytest rc=$?"; tail -8 $OUT/e_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 300 python bench.py --steps 40 --warmup 8 > $OUT/bench_1_all_reduce.json 2> $OUT/bench_1.err; cut -c1-400 $OUT/bench_1_all_reduce.json
timeout 300 python bench.py --steps 40 --warmup 8 --mode single > $OUT/bench_1_single.json 2>> $OUT/bench_1.err; cut -c1-200 $OUT/bench_1_single.json
timeout 300 python bench.py --steps 20 --warmup 5 --impl reference > $OUT/bench_1_ref_all_reduce.json 2>> $OUT/bench_1.err; cut -c1-200 $OUT/bench_1_ref_all_reduce.json
