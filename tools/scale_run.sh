#!/bin/bash
# Scaling evidence on one multi-GPU box: bench.py at N = 1, 2, 4, 8 and the fused-path rooflines at N = 8 (and 2).
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "gpus: $NG" > gpurun_out/scale.log
for N in 1 2 4 8; do
  [ $N -le $NG ] || continue
  if [ $N -eq 1 ]; then
    timeout 200 python bench.py --gpus 1 --steps 4 --warmup 3 --cpu-baseline-fits 0 2>/dev/null | tail -1 >> gpurun_out/scale.log
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 4 --warmup 3 --cpu-baseline-fits 0 2>/dev/null | tail -1 >> gpurun_out/scale.log
  fi
done
for N in 2 8; do
  [ $N -le $NG ] || continue
  echo "fused_paths N=$N" >> gpurun_out/scale.log
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+N)) bench/fused_paths.py --reps 4 2>/dev/null | grep '^{' >> gpurun_out/scale.log
done
echo "dist check N=$NG" >> gpurun_out/scale.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29799 tests/dist_fused_check.py 2>&1 | grep -E "^\[0\]|ALL OK|FAILED" >> gpurun_out/scale.log
cut -c1-420 gpurun_out/scale.log
