python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "attn or attention or flash" > gpurun_out/pytest_r02m_attn.log 2>&1; echo attn rc=$?; tail -3 gpurun_out/pytest_r02m_attn.log
python scripts/bench_attn.py > gpurun_out/bench_attn_r02m.txt 2>&1; cat gpurun_out/bench_attn_r02m.txt
python scripts/bench_configs.py --exo > gpurun_out/exo_r02m_n1.json 2> gpurun_out/exo_r02m_n1.err; echo exo rc=$?; tail -1 gpurun_out/exo_r02m_n1.json
python -m pytest tests/test_live_reference_gpu.py -m gpu -q -x -k "cfg3 or cfg5" > gpurun_out/pytest_r02m_live.log 2>&1; echo live rc=$?; tail -3 gpurun_out/pytest_r02m_live.log
