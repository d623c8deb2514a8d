set -x
python bench.py --steps 3 --warmup 3 --no-cpu --no-eager --sustain-s 0 > gpurun_out/bench_r02o_short.json 2> gpurun_out/bench_r02o_short.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 1400 --csv --log-file gpurun_out/launches_r02o.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-eager --sustain-s 0 > gpurun_out/ncu_launches_r02o.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:gemm_tc2 -s 300 -c 24 -o gpurun_out/ncu_full_r02o_gemm python bench.py --steps 3 --warmup 3 --no-cpu --no-eager --sustain-s 0 > gpurun_out/ncu_full_r02o.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:attn_flash -c 2 -o gpurun_out/ncu_full_r02o_flash python scripts/bench_attn.py 1024 > gpurun_out/ncu_flash_r02o.log 2>&1; echo rc=$?
ls -la gpurun_out/*.ncu-rep
