set -o pipefail
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
python tools/run_comm_once.py --mode zero > gpurun_out/rc_zero.log 2>&1 && $NCU -k regex:"zero_reduce_adamw|embed_fwd_sharded|zero_wait|gemm2_kernel<256, false, (false|true), false, true>" -c 9 -f -o gpurun_out/r2_comm_zero python tools/run_comm_once.py --mode zero > gpurun_out/rc_zero_ncu.log 2>&1; echo "zero ncu rc=$?"
python tools/run_comm_once.py --mode ddp > gpurun_out/rc_ddp.log 2>&1 && $NCU -k regex:"dp_reduce_adamw|dp_wait_done" -c 5 -f -o gpurun_out/r2_comm_ddp python tools/run_comm_once.py --mode ddp > gpurun_out/rc_ddp_ncu.log 2>&1; echo "ddp ncu rc=$?"
python tools/profile_decode.py > gpurun_out/pd_plain.log 2>&1 && $NCU -k regex:"gemm_skinny|attn_decode|sample_kernel" -c 9 -f -o gpurun_out/r2_decode python tools/profile_decode.py > gpurun_out/pd_ncu2.log 2>&1; echo "decode ncu rc=$?"
python tools/profile_step.py > gpurun_out/ps_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ps_ncu.log 2>&1; echo "launches rc=$?"
tail -1 gpurun_out/rc_zero.log gpurun_out/rc_ddp.log gpurun_out/ps_plain.log; ls -la gpurun_out/*.ncu-rep
