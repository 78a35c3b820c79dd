#!/bin/bash
# Multi-GPU session for the opt-in kernels / modes (run with  tools/g --gpus N --timeout 900 -- 'bash tools/gpu_session_multi.sh'):
# correctness first (dist check incl. the EXPERIMENTAL section), then timings with --experimental.
mkdir -p gpurun_out/session
NG=$(nvidia-smi -L | wc -l)
O=gpurun_out/session/multi_$NG.log
echo "gpus: $NG" > $O
CML_EXPERIMENTAL=1 CML_WAIT_TIMEOUT_S=15 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 \
  --master-port 29790 tests/dist_fused_check.py 2>&1 | grep -E "^\[0\]|ALL OK|FAILED|rror" | cut -c1-220 >> $O
CML_WAIT_TIMEOUT_S=15 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29791 \
  bench/fused_paths.py --reps 4 --experimental 2>/dev/null | grep '^{' >> $O
cut -c1-420 $O
