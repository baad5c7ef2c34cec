for v in "-DGEMM_PAIR=0" "-DGEMM_PAIR=1"; do
  echo "== $v"; F5E_NVCC_FLAGS="$v" timeout 200 python tools/microbench.py --only gemm 2>&1 | grep -E "qkv|ff1|ff2|gemm_out"
done
