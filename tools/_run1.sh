python -m pytest tests/test_round2_gpu.py tests/test_kernels_gpu.py tests/test_model_gpu.py -m gpu -x -q 2>&1 | tail -4
python tools/bench_generate.py --impl ours > gpurun_out/r2_gen_split.json 2> gpurun_out/r2_gen_split.err; cut -c1-260 gpurun_out/r2_gen_split.json
B200_DECODE_SPLITS=1 python tools/bench_generate.py --impl ours > gpurun_out/r2_gen_split1.json 2>&1; cut -c1-260 gpurun_out/r2_gen_split1.json
B200_DECODE_SPLITS=4 python tools/bench_generate.py --impl ours > gpurun_out/r2_gen_split4.json 2>&1; cut -c1-260 gpurun_out/r2_gen_split4.json
B200_SKINNY_ANY_SPLIT=1 python tools/bench_generate.py --impl ours > gpurun_out/r2_gen_any.json 2> gpurun_out/r2_gen_any.err; cut -c1-260 gpurun_out/r2_gen_any.json; tail -2 gpurun_out/r2_gen_any.err
B200_SKINNY_ANY_SPLIT=1 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "skinny or generate" 2>&1 | tail -2
