N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29800
for cfg in "X=1" "NCCL_ALGO=NVLS" "NCCL_ALGO=Tree" "NCCL_PROTO=Simple" "NCCL_PROTO=LL128" "NCCL_ALGO=Ring NCCL_PROTO=Simple NCCL_MIN_CTAS=32"; do
  P=$((P+1)); env $cfg timeout 120 $TR --master-port $P tools/nccl_probe.py 2>&1 | grep "MB:" 
done
P=$((P+1)); NCCL_DEBUG=INFO timeout 120 $TR --master-port $P tools/nccl_probe.py 2>&1 | grep -i "nvls\|channels\|algo" | head -12
