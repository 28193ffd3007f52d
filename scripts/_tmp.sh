set -u
mkdir -p gpurun_out
for c in 65536 32768 16384 65536; do
UVA_CHUNK_ROWS=$c timeout 600 python bench.py --steps 2 --warmup 2 --no-extras --no-cpu-baseline > gpurun_out/r03b_c$c.json 2> gpurun_out/r03b_c$c.err; python -c "
import json; d=json.load(open('gpurun_out/r03b_c$c.json')); print('chunk $c', d['value'], d['step_frac_of_burst_peak'], d['clocks']['sm_mhz'])"; tail -n 2 gpurun_out/r03b_c$c.err
done
