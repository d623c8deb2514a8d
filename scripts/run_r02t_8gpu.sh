TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541"
run() { name=$1; shift; timeout 240 $TR "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; }
run bench_r02t_n8 bench.py --gpus 8 --steps 30 --no-eager --no-cpu --sustain-s 0
run bench_r02t_n8_repl bench.py --gpus 8 --steps 30 --no-eager --no-cpu --sustain-s 0 --replicated-optimizer
run bench_r02t_n8_bf16 bench.py --gpus 8 --steps 30 --no-eager --no-cpu --sustain-s 0 --grad-comm bf16
run exo_r02t_n8 scripts/bench_configs.py --exo
for f in gpurun_out/bench_r02t_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith("{")][-1])
    print(sys.argv[1], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
grep "^{" gpurun_out/exo_r02t_n8.json | cut -c1-330
tail -3 gpurun_out/bench_r02t_n8.err
