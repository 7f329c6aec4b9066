python tools/profile_decode.py > gpurun_out/pd_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_decode.csv python tools/profile_decode.py > gpurun_out/pd_ncu.log 2>&1
tail -2 gpurun_out/pd_plain.log
