python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --model 1.3b --steps 10 --warmup 3 > gpurun_out/r2_b1_1p3b.json 2> gpurun_out/r2_b1_1p3b.err; echo "1.3b ours rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_1p3b.json | head -1)"
python bench.py --model 1.3b --steps 10 --warmup 3 --impl reference > gpurun_out/r2_b1_1p3b_ref.json 2> gpurun_out/r2_b1_1p3b_ref.err; echo "1.3b ref rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_1p3b_ref.json | head -1)"
B200_DETERMINISTIC=1 python bench.py --steps 20 --warmup 5 --skip-e2e > gpurun_out/r2_b1_det.json 2> gpurun_out/r2_b1_det.err; echo "deterministic rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_det.json | head -1)"
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_b1_now.json 2> gpurun_out/r2_b1_now.err; echo "now rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_b1_now.json | head -1) launches $(grep -o '"gpu_launches_per_step": [0-9]*' gpurun_out/r2_b1_now.json)"
python tools/bench_generate.py --impl ours > gpurun_out/r2_gen_ours.json 2> gpurun_out/r2_gen_ours.err; cut -c1-200 gpurun_out/r2_gen_ours.json
timeout 600 python tools/bench_generate.py --impl reference > gpurun_out/r2_gen_ref.json 2> gpurun_out/r2_gen_ref.err; cut -c1-200 gpurun_out/r2_gen_ref.json
timeout 400 python tools/bench_generate.py --impl reference --ref-autocast > gpurun_out/r2_gen_ref_bf16.json 2> gpurun_out/r2_gen_ref_bf16.err; cut -c1-200 gpurun_out/r2_gen_ref_bf16.json
