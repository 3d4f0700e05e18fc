for v in "" "$@"; do
  if [ -z "$v" ]; then unset HAVAC_CUDA_LIB; else export HAVAC_CUDA_LIB=tools/variants/libhavac_cuda_$v.so; fi
  echo "== variant ${v:-base}"
  python tools/quick_bench.py 1e8 35804 3 2>&1 | tail -2
done
