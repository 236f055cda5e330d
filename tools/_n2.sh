TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29901 tools/nvls_check.py 2>&1 | grep -v "^\[W\|^W1\|Warning" | tail -15
P=29910
run() { name=$1; shift; P=$((P+1)); env "$@" timeout 300 $TR --master-port $P bench.py --gpus 2 --no-cpu-baseline > gpurun_out/n2_$name.json 2> gpurun_out/n2_$name.err; echo "$name: $(python -c "import json,sys; d=json.loads(open('gpurun_out/n2_$name.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'])" 2>&1 | tail -1)"; }
python bench.py --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n1:', d['ms_per_step'])"
run nvls B2G_DP_NVLS=1
run nccl B2G_DP_NVLS=0
tail -5 gpurun_out/n2_nvls.err
