set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02n_tests.log 2>&1; echo "rc=$?"; tail -n 3 gpurun_out/r02n_tests.log
for f in 1 0; do
UVA_FUSED_PRENET=$f timeout 600 python bench.py --workload latency > gpurun_out/r02n_lat_f$f.json 2> gpurun_out/r02n_lat_f$f.err; python -c "
import json; d=json.load(open('gpurun_out/r02n_lat_f$f.json'))['detail']; print('fused=$f', {k:(round(v['p50_ms'],3), round(v['p50_ms_device_noise'],3)) for k,v in d.items()})"; tail -n 2 gpurun_out/r02n_lat_f$f.err
UVA_FUSED_PRENET=$f timeout 600 python bench.py --workload train --steps 20 --warmup 5 > gpurun_out/r02n_train_f$f.json 2> gpurun_out/r02n_train_f$f.err; python -c "
import json; d=json.load(open('gpurun_out/r02n_train_f$f.json')); print('fused=$f train', d['ms_per_step'], d['detail']['p50_ms'], d['detail']['max_ms'])"; tail -n 2 gpurun_out/r02n_train_f$f.err
done
