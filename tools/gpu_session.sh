#!/bin/bash
# One GPU box session that produces everything the docs still lack (run through tools/g, single GPU):
#   tools/g --timeout 1500 -- 'bash tools/gpu_session.sh'
# Every step runs under its own timeout; outputs land in gpurun_out/session/.
mkdir -p gpurun_out/session
O=gpurun_out/session
echo "== full GPU suite";           timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 | tee $O/pytest_gpu.txt
echo "== experimental GEMM-RS";     CML_EXPERIMENTAL=1 CML_WAIT_TIMEOUT_S=10 timeout 240 python -m pytest tests/test_gemm_rs_gpu.py tests/test_gemm_gpu.py -q -k "gemm_rs or grouped or single_rank or emulated" 2>&1 | tail -5 | tee $O/pytest_gemm_rs.txt
echo "== bench.py";                 timeout 300 python bench.py --steps 4 --warmup 3 2>/dev/null | tail -1 | tee $O/bench.json | cut -c1-300
echo "== pipeline bench";           timeout 400 python bench/pipeline_bench.py 2>/dev/null | tail -1 | tee $O/pipeline.json | cut -c1-600
echo "== stability bench";          timeout 300 python bench/stability_bench.py 2>&1 | tail -1 | tee $O/stability.json | cut -c1-600
echo "== pan-cancer bench";         timeout 600 python bench/pancancer_bench.py 2>&1 | tail -1 | tee $O/pancancer.json | cut -c1-600
echo "== stage bench";              timeout 300 python bench/stage_bench.py 2>&1 | grep "^{" | tee $O/stage_bench.jsonl | cut -c1-250
echo "== gemm bench";               timeout 200 python bench/gemm_bench.py 2>/dev/null | tee $O/gemm_bench.jsonl | cut -c1-250
echo "== lasso phases";             timeout 200 python tools/prof_lasso_phases.py 2>&1 | tail -3 | tee $O/lasso_phases.jsonl | cut -c1-300
echo "== ncu: CTA-pair GEMM (after the plain run above exited 0)"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gemm_tn_pair_kernel --launch-skip 1 --launch-count 1 \
  -o $O/prof_gemm_pair -f python tools/prof_kernels.py gemm 2>&1 | tail -2
