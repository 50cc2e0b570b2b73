# ncu --set full capture of the three fused render kernels (one launch each) on the short bench command
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-reference"
$CMD > gpurun_out/plain_full.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_ -s 9 -c 3 -f -o gpurun_out/prof_r2b $CMD > gpurun_out/ncu_full_r2b.log 2>&1
echo full_rc=$?
