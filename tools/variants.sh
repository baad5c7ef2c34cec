# developer tool: time one microbench group under several compile-time variants of the library
#   bash tools/variants.sh attn "-DATT_DEFER_OWAIT=0" "-DATT_DEFER_OWAIT=1"
grp=$1; shift
for v in "$@"; do
  echo "== $v"; F5E_NVCC_FLAGS="$v" timeout 300 python tools/microbench.py --only "$grp" 2>&1 | grep -vE "sdpa|cublas"
done
