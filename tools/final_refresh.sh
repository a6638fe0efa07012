#!/bin/bash
# Round-end refresh of profiles/: GPU tests, ncu captures of the hot kernel (-mr 2 / 3), bench.py both arms, launch list.
mkdir -p gpurun_out
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; tail -2 gpurun_out/final_pytest.log
for K in 2 3; do
  EXTRA=""; [ $K -eq 3 ] && EXTRA="--max-error 3"
  ncu --set full --clock-control none --import-source on -k regex:match_seed -c 1 -f -o gpurun_out/r01_seed_K$K python bench.py --steps 1 --warmup 3 --no-cpu $EXTRA > gpurun_out/ncu_K$K.log 2>&1
  python tools/ncu_summary.py gpurun_out/r01_seed_K$K.ncu-rep "python bench.py --steps 1 --warmup 3 --no-cpu $EXTRA" > gpurun_out/K$K.txt && cp gpurun_out/K$K.txt profiles/r01_ncu_match_seed_K$K.txt
done
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; tail -c 600 gpurun_out/final_bench.json
python bench.py --max-error 3 --no-cpu > gpurun_out/final_bench_mr3.json 2>/dev/null
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; cat gpurun_out/final_bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/final_ncu_launch.log 2>&1
