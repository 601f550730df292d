#!/bin/bash
# N-GPU exchange variants of the default workload, one JSON line each (run under gpurun --gpus N).
# usage: tools/scale_experiment.sh N outdir
N=${1:-8}; OUT=${2:-gpurun_out}; PORT=29600
run() {  # name, env, extra args
  local name=$1; shift; local envs=$1; shift
  PORT=$((PORT + 1))
  env $envs python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 8 --warmup 3 --zbuf-images 0 "$@" > $OUT/r02_scale_n${N}_${name}.log 2> $OUT/r02_scale_n${N}_${name}.err
  python - "$OUT/r02_scale_n${N}_${name}.log" "$name" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    e = d.get("e2e") or {}
    print(f"{sys.argv[2]:>22s}: value {d['value']/1e9:7.3f} G q/s  {d['ms_per_step']:6.3f} ms/pass  e2e {e.get('value', 0)/1e9:6.3f} G q/s  kernels {d['kernels_ms_per_step']}  verified {d.get('exchange_verified')}")
except Exception as ex:
    print(sys.argv[2], "FAILED", ex)
PY
}
run allgather "A=1"
run allgather_ctas8 "NCCL_MAX_CTAS=8" --no-e2e
run allgather_ctas16 "NCCL_MAX_CTAS=16" --no-e2e
run gather_root "A=1" --exchange gather --no-e2e
run p2p "A=1" --exchange p2p --no-e2e
run allgather_f32 "A=1" --exchange-wire f32 --no-e2e
PORT=$((PORT + 1))
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --workload zbuf450 \
    > $OUT/r02_zbuf450_n${N}.log 2> $OUT/r02_zbuf450_n${N}.err
python - "$OUT/r02_zbuf450_n${N}.log" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("zbuf450:", d["value"], "images/s", d["config"]["images_per_rank"], "images on rank 0;", d["roofline"]["frac"], "of the atomic sweep")
except Exception as ex:
    print("zbuf450 FAILED", ex)
PY
