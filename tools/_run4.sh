N=${N:-4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 400 $TR --master-port $((29600 + RANDOM % 300)) "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/$name.json | head -1) $(grep -o '"passed": [a-z]*' gpurun_out/$name.json)"; }
run r2_b${N}_ddp_u bench.py --gpus $N --steps 20 --warmup 5
run r2_b${N}_zero_u bench.py --gpus $N --steps 20 --warmup 5 --dp-mode zero
python bench.py --steps 20 --warmup 5 --skip-e2e > gpurun_out/r2_b1_plain.json 2>/dev/null; echo "plain 1 gpu $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_plain.json | head -1)"
python bench.py --steps 20 --warmup 5 --skip-e2e --force-dp > gpurun_out/r2_b1_forcedp.json 2>gpurun_out/r2_b1_forcedp.err; echo "force-dp 1 gpu $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_forcedp.json | head -1)"; tail -3 gpurun_out/r2_b1_forcedp.err
python bench.py --steps 20 --warmup 5 --skip-e2e --force-dp --dp-mode zero > gpurun_out/r2_b1_forcezero.json 2>gpurun_out/r2_b1_forcezero.err; echo "force-zero 1 gpu $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_forcezero.json | head -1)"; tail -3 gpurun_out/r2_b1_forcezero.err
