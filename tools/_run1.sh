python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_round2_gpu.py -m gpu -x -q -k "decode or generate" 2>&1 | tail -3
python tools/bench_attn_decode.py 2>&1 | tail -5 | cut -c1-200
python tools/bench_generate.py --impl ours > gpurun_out/r2_gen_v5.json 2> gpurun_out/r2_gen_v5.err; cut -c1-260 gpurun_out/r2_gen_v5.json
