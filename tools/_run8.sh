N=${N:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 600 $TR --master-port $((29600 + RANDOM % 300)) "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/$name.json | head -1) $(grep -o '"value": [0-9.]*' gpurun_out/$name.json | head -1) $(grep -o '"passed": [a-z]*' gpurun_out/$name.json)"; }
timeout 300 $TR --master-port 29901 tools/bench_dp_kernel.py --model gpt2-small > gpurun_out/r2_dpk${N}_small.jsonl 2> gpurun_out/r2_dpk${N}_small.err; echo "dpk small rc=$?"; cat gpurun_out/r2_dpk${N}_small.jsonl | cut -c1-420
timeout 300 $TR --master-port 29902 tools/bench_dp_kernel.py --model 1.3b > gpurun_out/r2_dpk${N}_1p3b.jsonl 2> gpurun_out/r2_dpk${N}_1p3b.err; echo "dpk 1.3b rc=$?"; cat gpurun_out/r2_dpk${N}_1p3b.jsonl | cut -c1-420
run r2_b${N}_ddp_u4 bench.py --gpus $N --steps 20 --warmup 5
B200_DP_UNROLL=1 run r2_b${N}_ddp_u1 bench.py --gpus $N --steps 20 --warmup 5 --no-dp-check --skip-e2e
run r2_b${N}_ddp_u4b bench.py --gpus $N --steps 20 --warmup 5 --no-dp-check --skip-e2e
run r2_b${N}_1p3b_zero bench.py --gpus $N --steps 10 --warmup 3 --model 1.3b --dp-mode zero
run r2_b${N}_1p3b_ref bench.py --impl reference --gpus $N --steps 10 --warmup 3 --model 1.3b
