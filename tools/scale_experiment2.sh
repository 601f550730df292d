#!/bin/bash
# Second set of N-GPU exchange variants (run under gpurun --gpus N): NCCL CTA caps, chunked copy-engine pushes, hybrid.
N=${1:-8}; OUT=${2:-gpurun_out}; PORT=29700
run() {
  local name=$1; shift; local envs=$1; shift
  PORT=$((PORT + 1))
  env $envs python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 8 --warmup 3 --zbuf-images 0 --no-e2e "$@" > $OUT/r02_scale_n${N}_${name}.log 2> $OUT/r02_scale_n${N}_${name}.err
  python - "$OUT/r02_scale_n${N}_${name}.log" "$name" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    k = d['kernels_ms_per_step']
    print(f"{sys.argv[2]:>24s}: value {d['value']/1e9:7.3f} G q/s  {d['ms_per_step']:6.3f} ms/pass  kernels {k['K_NN_KEYSORT']:.2f}+{k['K_NN_SEARCH']:.2f}+{k['K_NN_FAR']:.2f}+{k['K_ASSOC_RECORD']:.2f}  verified {d.get('exchange_verified')}")
except Exception as ex:
    print(sys.argv[2], "FAILED", ex)
PY
}
run ctas24 "NCCL_MAX_CTAS=24"
run p2p_chunks4 "A=1" --exchange p2p --p2p-chunks 4
run hybrid45_ctas16 "NCCL_MAX_CTAS=16" --exchange hybrid --hybrid-nccl-frac 0.45
run hybrid55_ctas24 "NCCL_MAX_CTAS=24" --exchange hybrid --hybrid-nccl-frac 0.55
run hybrid35_ctas8 "NCCL_MAX_CTAS=8" --exchange hybrid --hybrid-nccl-frac 0.30
