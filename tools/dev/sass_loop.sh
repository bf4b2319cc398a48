# Compiles plc_sweep.cu with the given flags to a cubin and prints k_sweep's register count, total SASS instructions and
# the instruction count of the march loop (between the two VOTE.ANY of the loop).  usage: sass_loop.sh NAME [FLAGS...]
name=$1; shift
cd "$(dirname "$0")/../../pointcloud_processor_b200/csrc"
mkdir -p _obj_var
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -I../../include -I. "$@" -cubin -Xptxas -v plc_sweep.cu -o _obj_var/$name.cubin 2>&1 | grep -A3 "k_sweepILb1" | grep -o "Used [0-9]* registers.*\|[0-9]* bytes spill stores"
cuobjdump -sass _obj_var/$name.cubin | awk '/Function : .*k_sweepILb1/{p=1} /Function : /{if(!/k_sweepILb1/)p=0} p' > _obj_var/$name.sass
grep -c "^ *\/\*[0-9a-f]\{4\}\*\/" _obj_var/$name.sass
