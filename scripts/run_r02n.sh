python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "attn or attention or flash" > gpurun_out/pytest_r02n_attn.log 2>&1; echo attn rc=$?; tail -15 gpurun_out/pytest_r02n_attn.log
python scripts/bench_attn.py 256 512 1024 2048 4096 > gpurun_out/bench_attn_r02n.txt 2>&1; cat gpurun_out/bench_attn_r02n.txt
python scripts/bench_configs.py --exo > gpurun_out/exo_r02n_n1.json 2> gpurun_out/exo_r02n_n1.err; echo exo rc=$?; tail -1 gpurun_out/exo_r02n_n1.json | cut -c1-400
python -m pytest tests/test_live_reference_gpu.py -m gpu -q -x -k "cfg3 or cfg5" > gpurun_out/pytest_r02n_live.log 2>&1; echo live rc=$?; tail -3 gpurun_out/pytest_r02n_live.log
