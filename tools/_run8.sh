N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 400 $TR --master-port $((29600 + RANDOM % 300)) "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/$name.json | head -1) $(grep -o '"value": [0-9.]*' gpurun_out/$name.json | head -1) $(grep -o '"passed": [a-z]*' gpurun_out/$name.json)"; }
run r2_final_b8 bench.py --gpus 8 --steps 20 --warmup 5
python bench.py --steps 20 --warmup 5 --skip-e2e > gpurun_out/r2_final_b1_samebox.json 2>/dev/null; echo "1 gpu same box $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_final_b1_samebox.json | head -1)"
run r2_final_b8_1p3b_zero bench.py --gpus 8 --steps 10 --warmup 3 --model 1.3b --dp-mode zero --skip-e2e
