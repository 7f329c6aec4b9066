python -m pytest tests/test_model_gpu.py tests/test_round2_gpu.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 --skip-e2e > gpurun_out/r2_b1_ws.json 2> gpurun_out/r2_b1_ws.err; echo "wstream $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_ws.json | head -1)"; tail -2 gpurun_out/r2_b1_ws.err
B200_WGRAD_STREAM=0 python bench.py --steps 20 --warmup 5 --skip-e2e > gpurun_out/r2_b1_nows.json 2> gpurun_out/r2_b1_nows.err; echo "single stream $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_nows.json | head -1)"
python bench.py --steps 20 --warmup 5 --skip-e2e > gpurun_out/r2_b1_ws2.json 2>/dev/null; echo "wstream again $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_ws2.json | head -1)"
python bench.py --steps 10 --warmup 3 --skip-e2e --model 1.3b > gpurun_out/r2_b1_ws_1p3b.json 2>/dev/null; echo "1.3b wstream $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_ws_1p3b.json | head -1)"
B200_WGRAD_STREAM=0 python bench.py --steps 10 --warmup 3 --skip-e2e --model 1.3b > gpurun_out/r2_b1_nows_1p3b.json 2>/dev/null; echo "1.3b single $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_nows_1p3b.json | head -1)"
