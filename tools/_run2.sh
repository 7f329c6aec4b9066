TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 500 $TR --master-port $((29600 + RANDOM % 300)) "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/$name.json | head -1) $(grep -o '"value": [0-9.]*' gpurun_out/$name.json | head -1)"; }
run r2_ref2_stock bench.py --impl reference --gpus 2 --steps 20 --warmup 5
REF_BROADCAST_BUFFERS=0 run r2_ref2_nobb bench.py --impl reference --gpus 2 --steps 20 --warmup 5
TORCH_COMPILE=0 run r2_ref2_eager bench.py --impl reference --gpus 2 --steps 20 --warmup 5
run r2_ours2 bench.py --gpus 2 --steps 20 --warmup 5
