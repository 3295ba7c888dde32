for v in "$@"; do
  echo "== $v"
  MULTIGRAD_B200_LIB=build/variants/libmgb200_$v.so python benchmarks/linlink_bench.py --halos 2e7 --features 64 2>&1 | grep kernel | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['frac_of_measured_copy_bw'])"
done
