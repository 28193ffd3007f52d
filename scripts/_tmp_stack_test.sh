set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_headline.py -x -q -k "layer_chain" > gpurun_out/r02f_stack.log 2>&1; echo "rc=$?"; tail -n 8 gpurun_out/r02f_stack.log
for g in 16 24 32 0; do
UVA_STACK_GROUP=$g timeout 600 python bench.py --steps 2 --warmup 2 --no-extras --no-cpu-baseline > gpurun_out/r02f_bench_g$g.json 2> gpurun_out/r02f_bench_g$g.err; echo "group $g rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r02f_bench_g$g.json')); print(d['value'], d['ms_per_step'], d['step_frac_of_burst_peak'], d['clocks'], d['gpu_launches'])"; tail -n 2 gpurun_out/r02f_bench_g$g.err
done
