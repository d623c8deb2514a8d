for v in "" "EXG_NO_GEMM_PAIR=1" "EXG_GEMM_NCTA=2" "EXG_GEMM_NCTA=1"; do
  env $v python bench.py --no-eager --no-cpu --sustain-s 0 --steps 40 > gpurun_out/bench_r02p.tmp 2>/dev/null
  python - "$v" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/bench_r02p.tmp").read().strip().splitlines()[-1])
print(f"{sys.argv[1] or 'default':24s} {d['value']:9.1f} clips/s {d['ms_per_step']:.4f} ms  gemm {d['breakdown']['gemm_tc']['ms']:.3f} ms {d['breakdown']['gemm_tc']['tflops']:.0f} TF/s  clocks {d['clocks']['sm_mhz']}")
PY
done
