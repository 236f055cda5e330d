run() { env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('$*', round(d['value']), round(d['ms_per_step'],3))"; PORT=$((PORT+1)); }
PORT=29570
run B2G_DP_SM_RESERVE=0
run B2G_DP_SM_RESERVE=8
run B2G_DP_SM_RESERVE=16
run B2G_DP_SM_RESERVE=32
run B2G_DP_SM_RESERVE=16 B2G_DP_SM_RESERVE_BWD=1
run B2G_DP_SM_RESERVE=16 NCCL_MAX_NCTAS=16
run B2G_DP_SM_RESERVE=8 NCCL_MAX_NCTAS=8
run B2G_DP_SM_RESERVE=0 NCCL_MAX_NCTAS=8
run B2G_DP_SM_RESERVE=0
