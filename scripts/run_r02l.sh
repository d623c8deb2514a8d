python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "attn or attention" > gpurun_out/pytest_r02l_attn.log 2>&1; echo attn rc=$?; tail -3 gpurun_out/pytest_r02l_attn.log
python scripts/bench_attn.py 32 64 96 128 > gpurun_out/bench_attn_r02l.txt 2>&1; cat gpurun_out/bench_attn_r02l.txt
python bench.py --no-eager --no-cpu --steps 30 > gpurun_out/bench_r02l_n1.json 2> gpurun_out/bench_r02l_n1.err; echo n1 rc=$?
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_r02l_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], {k:(v["ms"],v["launches"]) for k,v in d["breakdown"].items()})
PY
python scripts/profile_gemm_shapes.py > gpurun_out/gemm_shapes_tan_r02l.txt 2>&1; cat gpurun_out/gemm_shapes_tan_r02l.txt
python scripts/profile_gemm_shapes.py --exo > gpurun_out/gemm_shapes_exo_r02l.txt 2>&1; cat gpurun_out/gemm_shapes_exo_r02l.txt
python -m pytest tests/test_model_gpu.py tests/test_train_step_gpu.py tests/test_live_reference_gpu.py -m gpu -q -x > gpurun_out/pytest_r02l_model.log 2>&1; echo model rc=$?; tail -3 gpurun_out/pytest_r02l_model.log
