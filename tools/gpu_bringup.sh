#!/bin/bash
# Runs each GPU kernel test group in its own process under a timeout, so a hang or fault in one
# kernel does not hide the results of the others.  Output: gpurun_out/bringup/*.log + summary.txt
cd "$(dirname "$0")/.."
OUT=gpurun_out/bringup
mkdir -p $OUT
: > $OUT/summary.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/gpu.txt 2>&1
TESTS=${@:-"test_gemm_tn_bf16 test_gemm_epilogues test_gemm_nt_dgrad test_gemm_tt_wgrad test_gemm_stats test_layernorm_fwd_bwd test_embed test_adamw_matches_torch test_colsum_cast test_flash_attention_fwd_bwd test_decode_attention test_sampler_distribution"}
for t in $TESTS; do
  timeout 420 python -m pytest tests/test_kernels_gpu.py -q -k "$t" -p no:cacheprovider > $OUT/$t.log 2>&1
  rc=$?
  echo "$t rc=$rc $(tail -1 $OUT/$t.log)" | tee -a $OUT/summary.txt
done
