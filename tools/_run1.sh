python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/r2_final_b1.json 2> gpurun_out/r2_final_b1.err; echo "bench rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_final_b1.json | head -1) $(grep -o '"value": [0-9.]*' gpurun_out/r2_final_b1.json | head -1)"
python bench.py --impl reference > gpurun_out/r2_final_b1_ref.json 2> gpurun_out/r2_final_b1_ref.err; echo "ref rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_final_b1_ref.json | head -1) $(grep -o '"value": [0-9.]*' gpurun_out/r2_final_b1_ref.json | head -1)"
