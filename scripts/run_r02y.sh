python bench.py --steps 3 --warmup 3 --no-cpu --no-eager --sustain-s 0 > gpurun_out/bench_r02y_short.json 2> gpurun_out/bench_r02y_short.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 1400 --csv --log-file gpurun_out/launches_r02y.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-eager --sustain-s 0 > gpurun_out/ncu_launches_r02y.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 20 -c 8 -o gpurun_out/ncu_full_r02y_attn python bench.py --steps 2 --warmup 3 --no-cpu --no-eager --sustain-s 0 --no-graph > gpurun_out/ncu_attn_r02y.log 2>&1; echo rc=$?
ls -la gpurun_out/*r02y*
