#!/bin/bash
# Reads the cost of the data-parallel collectives off the step time: bench.py at N ranks with collectives dropped (DV3_DP_SKIP, results
# are wrong then) or NCCL steered differently.  usage: tools/dp_probe.sh N [size] [short]
N=${1:-8}; SZ=${2:-S}
run() {
  echo -n "$1: "
  env $2 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 200)) \
    bench.py --gpus $N --size $SZ --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys
l=[x for x in sys.stdin if x.startswith('{')]
b=json.loads(l[-1]) if l else None
print(round(b['value'],1), round(b['ms_per_step'],3)) if b else print('failed')"
}
run "base" "A=1"
if [ "$3" = "short" ]; then run "no collectives" "DV3_DP_SKIP=reduce,gather"; exit 0; fi
run "no collectives" "DV3_DP_SKIP=reduce,gather"
run "no all-reduce" "DV3_DP_SKIP=reduce"
run "no all-gather" "DV3_DP_SKIP=gather"
run "no mid bucket" "DV3_DP_MID_BUCKET=0"
run "NCCL_ALGO=Tree" "NCCL_ALGO=Tree"
run "NCCL_ALGO=Ring" "NCCL_ALGO=Ring"
run "NCCL_NVLS_ENABLE=0" "NCCL_NVLS_ENABLE=0"
run "NCCL_MAX_NCHANNELS=8" "NCCL_MAX_NCHANNELS=8"
