# A/B of k_sweep variants (make variant NAME=...): kernel_ms on cfg2 (3 runs each) and cfg3 (1 run).  usage: ab3.sh main v1 v2 ...
for v in "$@"; do
  if [ "$v" = main ]; then unset PLC_DEV_LIB; else export PLC_DEV_LIB=$PWD/pointcloud_processor_b200/lib/dev_$v.so; fi
  a=$(for i in 1 2 3; do python tools/counters.py cfg2 2>&1 | tail -1 | python -c "import sys,json; print(round(json.loads(sys.stdin.read())['kernel_ms'],4))"; done | tr '\n' ' ')
  b=$(python tools/counters.py cfg3 2>&1 | tail -1 | python -c "import sys,json; print(round(json.loads(sys.stdin.read())['kernel_ms'],3))")
  echo "$v cfg2: $a cfg3: $b"
done
