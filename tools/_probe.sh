SCT_TRACE=1 python tools/e2e_probe.py 1000000 2>&1 | grep -m2 "seeded"
SCT_TRACE=1 python tools/e2e_probe.py 1000000 seed_prefilter=0 2>&1 | grep -m1 "seeded"
ncu --set full --clock-control none --import-source on -k regex:match_seed_kernel -c 1 -o gpurun_out/r01_seed_pre python bench.py --no-cpu --steps 1 --warmup 0 > gpurun_out/ncu_seed.log 2>&1; tail -1 gpurun_out/ncu_seed.log
