# Round evidence in one gpurun call: GPU suite, both bench arms, launch list, --set full capture of k_sweep (cfg2).
# usage: bash tools/dev/r2_evidence.sh [tag]
tag=${1:-r2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${tag}_smi.txt 2>&1
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/${tag}_gpu_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_gpu_tests.txt
timeout 600 python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref_n1.json 2> gpurun_out/${tag}_bench_ref_n1.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-north-star --no-feeders > gpurun_out/${tag}_ncu_bench.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sweep -s 1 -c 1 -f -o gpurun_out/${tag}_k_sweep \
  python tools/counters.py cfg2 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "full capture rc=$?"
tail -3 gpurun_out/${tag}_gpu_tests.txt; cat gpurun_out/${tag}_bench_n1.json | cut -c1-1500
# cfg5 (4000 seeded poses, refinement as the full config gets it): full capture of k_sweep
PLC_DEV_REFINE=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sweep -s 1 -c 1 -f -o gpurun_out/${tag}_k_sweep_cfg5 \
  python tools/counters_subset.py cfg5 4000 > gpurun_out/${tag}_ncu_cfg5.log 2>&1; echo "cfg5 capture rc=$?"
