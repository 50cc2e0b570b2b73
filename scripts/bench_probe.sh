# the short bench command the ncu passes profile (environment recorded beside it)
env | sort > gpurun_out/env_$$.log; ulimit -a >> gpurun_out/env_$$.log; ls -la /proc/$$/fd >> gpurun_out/env_$$.log 2>&1
exec python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-reference
