mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 295$1$2 bench.py --gpus $1 --no-cpu-baseline $3 2>> gpurun_out/scale.err; }
run 8 0 "" > gpurun_out/r02_bench_n8.json
run 4 1 "" > gpurun_out/r02_bench_n4.json
run 2 2 "" > gpurun_out/r02_bench_n2.json
run 8 3 "--scaling weak" > gpurun_out/r02_bench_n8_weak.json
for f in n2 n4 n8 n8_weak; do python -c "
import json,sys; d=json.load(open('gpurun_out/r02_bench_$f.json')); print('$f', d['n_gpus'], d['scaling'], round(d['value']/1e9,3), 'G samples/s', round(d['rays_per_s']/1e9,2), 'G rays/s', round(d['ms_per_step'],2), 'ms e2e', round(d['e2e']['value']/1e9,3), d['e2e']['step_ms_rank0'], d['clocks'])"; done
tail -3 gpurun_out/scale.err
