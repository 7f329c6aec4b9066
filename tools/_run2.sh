TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for mode in zero ddp; do
timeout 300 $TR --master-port $((29700 + RANDOM % 200)) scripts/train_transformer.py preset=tiny synthetic_data=true dp_mode=$mode t_train_steps=30 t_eval_steps=10 t_eval_iters=2 save_every=20 t_out_path=/tmp/ck_$mode.pt log_jsonl=/tmp/log_$mode.jsonl > gpurun_out/r2_script_$mode.log 2>&1; echo "train script $mode rc=$?"; grep -h "Step \|finished\|Error\|error" gpurun_out/r2_script_$mode.log | head -6
python scripts/generate_text.py --model_path /tmp/ck_$mode.pt --input_ids 1,2,3,4 --max_new_tokens 12 2>&1 | tail -2
done
