N=$1; shift
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29700
run() { name=$1; shift; P=$((P+1)); env "$@" timeout 400 $TR --master-port $P bench.py --gpus $N --no-cpu-baseline $EXTRA > gpurun_out/n${N}_$name.json 2> gpurun_out/n${N}_$name.err; echo "n$N $name: $(python -c "import json,sys; d=json.loads(open('gpurun_out/n${N}_$name.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'])" 2>&1 | tail -1)"; }
python bench.py --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n1:', d['ms_per_step'])"
run old B2G_DP_G_CTAS=0 B2G_DP_EARLY_ALLREDUCE=1
run g4w2 B2G_DP_G_CTAS=4
run g6w2 B2G_DP_G_CTAS=6
run g8w2 B2G_DP_G_CTAS=8
EXTRA="--global-batch 512" run gb512_g4w2 B2G_DP_G_CTAS=4
P=$((P+1)); B2G_DP_G_CTAS=4 $TR --master-port $P tools/dp_timeline.py --graph 1 > gpurun_out/dp_tl_n${N}_g1_ctas4.log 2>&1; tail -3 gpurun_out/dp_tl_n${N}_g1_ctas4.log | cut -c1-1200
