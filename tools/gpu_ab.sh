#!/bin/bash
for i in 1 2; do
  for v in 1 0; do
    echo -n "ACTFUSE=$v  "; B200_GEMM_ACTFUSE=$v timeout 300 python tools/pp_model.py llama2-7b q4_0 512 8 2>&1 | tail -1
  done
done
for v in 1 0; do echo -n "ACTFUSE=$v  "; B200_GEMM_ACTFUSE=$v timeout 300 python tools/pp_model.py llama3-8b q4_K_M 512 8 2>&1 | tail -1; done
