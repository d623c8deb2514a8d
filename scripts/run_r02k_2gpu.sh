TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
python -m pytest tests/test_xneg_gpu.py tests/test_train_step_gpu.py -m gpu -q > gpurun_out/pytest_r02k_2gpu.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest_r02k_2gpu.log
python bench.py --no-eager --no-cpu --steps 30 > gpurun_out/bench_r02k_n1.json 2> gpurun_out/bench_r02k_n1.err; echo n1 rc=$?
$TR bench.py --gpus 2 --steps 30 --no-eager --no-cpu > gpurun_out/bench_r02k_n2.json 2> gpurun_out/bench_r02k_n2.err; echo n2 rc=$?
$TR bench.py --gpus 2 --steps 30 --no-eager --no-cpu --comm-chunks 1 --sustain-s 0 > gpurun_out/bench_r02k_n2_c1.json 2>> gpurun_out/bench_r02k_n2.err; echo c1 rc=$?
$TR bench.py --gpus 2 --steps 30 --no-eager --no-cpu --comm-chunks 8 --sustain-s 0 > gpurun_out/bench_r02k_n2_c8.json 2>> gpurun_out/bench_r02k_n2.err; echo c8 rc=$?
$TR bench.py --gpus 2 --steps 30 --no-eager --no-cpu --grad-comm bf16 --sustain-s 0 > gpurun_out/bench_r02k_n2_bf16.json 2>> gpurun_out/bench_r02k_n2.err; echo bf16 rc=$?
$TR bench.py --gpus 2 --steps 30 --no-eager --no-cpu --host-dtype bf16 --sustain-s 0 > gpurun_out/bench_r02k_n2_hbf16.json 2>> gpurun_out/bench_r02k_n2.err; echo hbf16 rc=$?
$TR bench.py --gpus 2 --steps 30 --no-eager --no-cpu --xneg --sustain-s 0 > gpurun_out/bench_r02k_n2_xneg.json 2>> gpurun_out/bench_r02k_n2.err; echo xneg rc=$?
$TR scripts/bench_configs.py --exo > gpurun_out/exo_r02k_n2.json 2> gpurun_out/exo_r02k_n2.err; echo exo rc=$?
python scripts/bench_configs.py --exo > gpurun_out/exo_r02k_n1.json 2> gpurun_out/exo_r02k_n1.err; echo exo1 rc=$?
for f in gpurun_out/bench_r02k_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["config"].get("host_placement"))
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
tail -3 gpurun_out/exo_r02k_n2.json gpurun_out/exo_r02k_n1.json
tail -5 gpurun_out/bench_r02k_n2.err
