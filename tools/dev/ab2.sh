# usage: bash tools/dev/ab2.sh <scene> <variant>...   (variant "main" = libplc.so)
scene=$1; shift
for v in "$@"; do
  for i in 1 2; do
    if [ "$v" = main ]; then python tools/counters.py $scene 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$scene $v', round(d['kernel_ms'],4), d['samples'], d['rays_visible'])";
    else PLC_DEV_LIB=$PWD/pointcloud_processor_b200/lib/dev_$v.so python tools/counters.py $scene 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$scene $v', round(d['kernel_ms'],4), d['samples'], d['rays_visible'])"; fi
  done
done
