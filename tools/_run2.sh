TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
DP_MODE=zero timeout 200 $TR --master-port 29912 tools/dp_trainer_check.py > gpurun_out/r2_fin_ztr.log 2>&1; echo "trainer zero accum=2 rc=$?"; grep -h "checkpoint vs\|after resume\|DP_TRAINER\|Error" gpurun_out/r2_fin_ztr.log | head -5
