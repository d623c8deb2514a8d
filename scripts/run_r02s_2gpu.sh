TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 500 python -m pytest tests/test_xneg_gpu.py -m gpu -q > gpurun_out/pytest_r02s_2gpu.log 2>&1; echo pytest rc=$?; tail -4 gpurun_out/pytest_r02s_2gpu.log
run() { name=$1; shift; timeout 200 $TR "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; }
run bench_r02s_n2 bench.py --gpus 2 --steps 30 --no-eager --no-cpu --sustain-s 0
run bench_r02s_n2_repl bench.py --gpus 2 --steps 30 --no-eager --no-cpu --sustain-s 0 --replicated-optimizer
run bench_r02s_n2_bf16 bench.py --gpus 2 --steps 30 --no-eager --no-cpu --sustain-s 0 --grad-comm bf16
timeout 200 python bench.py --steps 30 --no-eager --no-cpu --sustain-s 0 > gpurun_out/bench_r02s_n1.json 2> gpurun_out/bench_r02s_n1.err; echo n1 rc=$?
for f in gpurun_out/bench_r02s_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith("{")][-1])
    print(sys.argv[1], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
tail -5 gpurun_out/bench_r02s_n2.err
