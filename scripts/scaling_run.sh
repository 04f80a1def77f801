#!/bin/sh
# Multi-GPU measurements of one box: bench.py at N = 1,2,4,8 (weak scaling), config 4 at 1/2/4/8 and
# config 5 at 8 (strong scaling: the samples are split).  Output: gpurun_out/scale_*.json
OUT=${1:-gpurun_out}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/scale_bench_n1.json 2> $OUT/scale_bench_n1.err
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600 + n)) bench.py --gpus $n --steps 3 --warmup 3 > $OUT/scale_bench_n$n.json 2> $OUT/scale_bench_n$n.err
done
python scripts/render_config.py --config 4 --repeat 2 > $OUT/scale_config4_n1.json 2> $OUT/scale_config4_n1.err
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29700 + n)) scripts/render_config.py --config 4 --repeat 2 > $OUT/scale_config4_n$n.json 2> $OUT/scale_config4_n$n.err
done
$TR --nproc-per-node 8 --master-port 29800 scripts/render_config.py --config 5 > $OUT/scale_config5_n8.json 2> $OUT/scale_config5_n8.err
python scripts/render_config.py --config 3 > $OUT/scale_config3_n1.json 2> $OUT/scale_config3_n1.err
grep -h '"' $OUT/scale_*.json | cut -c1-400
