set -u
mkdir -p gpurun_out
for v in 0 1 0 1; do
UVA_QUAD=$v timeout 600 python bench.py --steps 2 --warmup 2 --no-extras --no-cpu-baseline > gpurun_out/r02y_bench_q$v.json 2> gpurun_out/r02y_bench_q$v.err; echo "quad $v rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r02y_bench_q$v.json')); print(d['value'], d['ms_per_step'], d['step_frac_of_burst_peak'], d['clocks'])"; tail -n 2 gpurun_out/r02y_bench_q$v.err
done
