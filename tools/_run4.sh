N=${N:-4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 400 $TR --master-port $((29600 + RANDOM % 300)) "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; cut -c1-230 gpurun_out/$name.json; grep -o '"dp_check".*' gpurun_out/$name.json | cut -c1-300; }
timeout 300 $TR --master-port 29911 tools/dp_check.py > gpurun_out/r2_dp${N}.log 2>&1; echo "dp_check ddp rc=$?"
grep -n "world\|mean|dW|\|Error\|DP_CHECK" gpurun_out/r2_dp${N}.log | head -5
run r2_b${N}_ddp bench.py --gpus $N --steps 20 --warmup 5
B200_DP_MULTIMEM=0 run r2_b${N}_ddp_p2p bench.py --gpus $N --steps 20 --warmup 5 --no-dp-check
B200_DP_SHARD_EMBED=0 run r2_b${N}_ddp_pushembed bench.py --gpus $N --steps 20 --warmup 5 --no-dp-check
