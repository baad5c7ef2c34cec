F5E_NVCC_FLAGS="" timeout 200 python tools/microbench.py --only attn 2>&1 | head -1
