#!/bin/bash
# Round-end refresh on ONE B200 box: GPU tests, the bench line (both arms), CLI end to end, the alignment at full size,
# ncu summaries of the final kernels.  Outputs land in gpurun_out/ and are copied into profiles/ by hand.
set -u
mkdir -p gpurun_out
(python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3) > gpurun_out/final_tests_alone.log; (python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/final_tests.log
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
python bench.py --impl reference > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err
python tools/cli_e2e.py 10000000 2 1 > gpurun_out/final_cli_10M.json 2> gpurun_out/final_cli_10M.err
python tools/bench_align.py --no-cpu --reads 10000000 --steps 3 --warmup 1 > gpurun_out/final_align_10M.json 2> gpurun_out/final_align_10M.err
python tools/bench_align.py > gpurun_out/final_align_200k.json 2> gpurun_out/final_align_200k.err
for k in 2 3; do
  ncu --set full --clock-control none --import-source on -k regex:match_seed_kernel -c 1 -f -o gpurun_out/final_seed_k$k \
      python bench.py --segments 1000000 --max-error $k --steps 1 --warmup 1 --no-cpu --no-secondary > gpurun_out/final_ncu_k$k.log 2>&1
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/final_launches_bench.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-secondary > gpurun_out/final_ncu_launches.log 2>&1
cat gpurun_out/final_tests_alone.log gpurun_out/final_tests.log
