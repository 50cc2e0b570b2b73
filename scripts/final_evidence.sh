# round-2 final evidence: GPU tests, bench line, ncu launch list and ncu --set full capture of the three fused kernels
set -x
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/gputests.log 2>&1; echo tests_rc=$?
tail -n 3 gpurun_out/gputests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo smoke_rc=$?
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo bench_rc=$?
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref_rc=$?
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-reference"
$CMD > gpurun_out/plain_r2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_r2.csv $CMD > gpurun_out/ncu_launch_r2.log 2>&1
echo launches_rc=$?
$CMD > gpurun_out/plain_r2b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_ -s 9 -c 3 -f -o gpurun_out/prof_r2 $CMD > gpurun_out/ncu_full_r2.log 2>&1
echo full_rc=$?
