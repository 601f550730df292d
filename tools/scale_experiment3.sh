#!/bin/bash
# Third set (run under gpurun --gpus N): more pieces per peer for the copy-engine exchange.
N=${1:-8}; OUT=${2:-gpurun_out}; PORT=29800
for C in ${CHUNKS:-8 16}; do
  PORT=$((PORT + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 8 --warmup 3 --zbuf-images 0 --no-e2e --exchange p2p --p2p-chunks $C > $OUT/r02_scale_n${N}_p2p_chunks$C.log 2> $OUT/r02_scale_n${N}_p2p_chunks$C.err
  python - "$OUT/r02_scale_n${N}_p2p_chunks$C.log" "p2p_chunks$C" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); k = d['kernels_ms_per_step']
    print(f"{sys.argv[2]:>16s}: value {d['value']/1e9:7.3f} G q/s  {d['ms_per_step']:6.3f} ms/pass  kernels {k['K_NN_KEYSORT']:.2f}+{k['K_NN_SEARCH']:.2f}+{k['K_NN_FAR']:.2f}+{k['K_ASSOC_RECORD']:.2f}  verified {d.get('exchange_verified')}")
except Exception as ex:
    print(sys.argv[2], "FAILED", ex)
PY
done
