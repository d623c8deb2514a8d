EXG_GEMM_DIRECT=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "gemm" 2>&1 | tail -3
S="qkv,out_proj,c_proj,du (dx,sim,dh =,dxn"
echo "--- default"; ONLY="$S" NOCUBLAS=1 timeout 120 python scripts/bench_gemm.py 2>&1 | grep tc2 | cut -c1-75
echo "--- direct"; EXG_GEMM_DIRECT=1 ONLY="$S" NOCUBLAS=1 timeout 120 python scripts/bench_gemm.py 2>&1 | grep tc2 | cut -c1-75
for v in "" "EXG_GEMM_DIRECT=1"; do env $v timeout 200 python bench.py --no-eager --no-cpu --sustain-s 0 --steps 40 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v', d['value'], d['ms_per_step'], 'gemm', d['breakdown']['gemm_tc']['ms'], d['e2e']['last_loss'])"; done
