timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "attn or attention or flash" > gpurun_out/pytest_r02w_attn.log 2>&1; echo attn rc=$?; tail -12 gpurun_out/pytest_r02w_attn.log
timeout 200 python scripts/bench_attn.py 256 512 1024 2048 4096 > gpurun_out/bench_attn_r02w.txt 2>&1; cat gpurun_out/bench_attn_r02w.txt
EXG_FLASH_FWD_V2=1 timeout 200 python scripts/bench_attn.py 1024 4096 2>&1 | tail -2
timeout 300 python -m pytest tests/test_live_reference_gpu.py -m gpu -q -x -k "cfg3 or cfg5" > gpurun_out/pytest_r02w_live.log 2>&1; echo live rc=$?; tail -3 gpurun_out/pytest_r02w_live.log
timeout 200 python scripts/bench_configs.py --exo > gpurun_out/exo_r02w_n1.json 2> gpurun_out/exo_r02w_n1.err; echo exo rc=$?; tail -1 gpurun_out/exo_r02w_n1.json | cut -c180-330
