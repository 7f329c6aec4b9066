python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -m gpu -x -q -k "skinny or generate" 2>&1 | tail -2
python tools/bench_generate.py --impl ours > gpurun_out/r2_gen_v3.json 2> gpurun_out/r2_gen_v3.err; cut -c1-260 gpurun_out/r2_gen_v3.json
python tools/bench_generate.py --impl ours --model gpt2-small > gpurun_out/r2_gen_v3_small.json 2>&1; cut -c1-260 gpurun_out/r2_gen_v3_small.json
