# 8-GPU pass: bench.py at N = 8, 4, 2 (weak scaling, cbox) and the C5 strong-scaling line at N = 8, 4, 2, 1
set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
  $TR --nproc-per-node $n --master-port $((29500 + n)) bench.py --gpus $n --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_n$n.json 2> gpurun_out/bench_n$n.err
done
rm -f gpurun_out/r01_c5_scale.jsonl
for n in 8 4 2; do
  $TR --nproc-per-node $n --master-port $((29600 + n)) tools/c5_scale.py --spp 64 --steps 2 >> gpurun_out/r01_c5_scale.jsonl 2> gpurun_out/c5_n$n.err
done
python tools/c5_scale.py --spp 64 --steps 2 >> gpurun_out/r01_c5_scale.jsonl 2> gpurun_out/c5_n1.err
tail -c 600 gpurun_out/r01_c5_scale.jsonl
