set -u
mkdir -p gpurun_out
for v in 0 1 0 1; do
UVA_ZIGZAG=$v timeout 600 python bench.py --steps 2 --warmup 2 --no-extras --no-cpu-baseline > gpurun_out/r02s_bench_z$v.json 2> gpurun_out/r02s_bench_z$v.err; echo "zigzag $v rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r02s_bench_z$v.json')); print(d['value'], d['ms_per_step'], d['step_frac_of_burst_peak'], d['clocks'])"; tail -n 2 gpurun_out/r02s_bench_z$v.err
done
timeout 900 python -m pytest tests/test_gpu_headline.py tests/test_gpu_infer.py -x -q > gpurun_out/r02s_tests.log 2>&1; echo "rc=$?"; tail -n 3 gpurun_out/r02s_tests.log
