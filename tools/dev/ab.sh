for v in "$@"; do
  for i in 1 2 3; do
    if [ "$v" = main ]; then python tools/counters.py cfg2 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', round(d['kernel_ms'],4))";
    else PLC_DEV_LIB=$PWD/pointcloud_processor_b200/lib/dev_$v.so python tools/counters.py cfg2 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', round(d['kernel_ms'],4))"; fi
  done
done
