TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 python tools/zero_check1.py > gpurun_out/r2_z1.log 2>&1; echo "zero_check1 rc=$?"; grep -v Warn gpurun_out/r2_z1.log | tail -8
DP_MODE=zero timeout 300 $TR --master-port 29601 tools/dp_check.py > gpurun_out/r2_dp2z.log 2>&1; echo "dp_check zero rc=$?"
grep -n "world 2\|mean|dW|\|Error\|error\|DP_CHECK" gpurun_out/r2_dp2z.log | head
timeout 300 $TR --master-port 29602 tools/dp_check.py > gpurun_out/r2_dp2.log 2>&1; echo "dp_check ddp rc=$?"
grep -n "world 2\|mean|dW|\|Error\|error\|DP_CHECK" gpurun_out/r2_dp2.log | head
timeout 300 $TR --master-port 29603 bench.py --gpus 2 --steps 20 --warmup 5 --dp-mode zero > gpurun_out/r2_b2z.json 2> gpurun_out/r2_b2z.err; echo "bench2 zero rc=$?"
cut -c1-250 gpurun_out/r2_b2z.json; grep -o '"arm_details".*' gpurun_out/r2_b2z.json | cut -c1-900; grep -i "symm\|error\|Traceback" gpurun_out/r2_b2z.err | head
