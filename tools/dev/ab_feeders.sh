# A/B of feeder variants: feeders_large kernel times per library.  usage: ab_feeders.sh main v1 v2 ...
for v in "$@"; do
  if [ "$v" = main ]; then unset PLC_DEV_LIB; else export PLC_DEV_LIB=$PWD/pointcloud_processor_b200/lib/dev_$v.so; fi
  python tools/dev/filter_large.py 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$v', ' '.join(f\"{k}={v['kernel_ms']:.4f}\" for k,v in d.items() if isinstance(v,dict)))"
done
