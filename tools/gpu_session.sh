#!/bin/bash
# One GPU box session that produces everything the docs still lack (run through tools/g, single GPU):
#   tools/g --timeout 1700 -- 'bash tools/gpu_session.sh'
# Every step runs under its own timeout; outputs land in gpurun_out/session/.  Order: correctness of what was validated in
# round 1, then the round-2 kernels (same cases as the CPU suite ran on the host execution model), then the benches (each
# records SM clocks per measurement), then one ncu capture per changed hot kernel - each only after its plain run exited 0.
mkdir -p gpurun_out/session
O=gpurun_out/session
echo "== full GPU suite";           timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 | tee $O/pytest_gpu.txt
echo "== round-2 kernels only";     timeout 300 python -m pytest tests/test_y_kernels_gpu.py -q -m gpu 2>&1 | tail -5 | tee $O/pytest_round2.txt
echo "== lasso determinism";        timeout 120 python tools/determinism_lasso.py 2>&1 | tail -6 | tee $O/determinism.txt
echo "== bench.py";                 timeout 400 python bench.py --steps 5 --warmup 3 2>$O/bench.err | tail -1 | tee $O/bench.json | cut -c1-400
echo "== bench.py reference arm";   timeout 300 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 | tee $O/bench_ref.json | cut -c1-300
echo "== lasso latency (team vs warp kernel)"
for M in 0 1; do CML_LASSO_TEAM=$M timeout 200 python tools/time_lasso_reps.py 2>&1 | tail -2 | sed "s/^/team=$M /" | tee -a $O/lasso_latency.txt; done
echo "== pipeline bench";           timeout 400 python bench/pipeline_bench.py 2>/dev/null | tail -1 | tee $O/pipeline.json | cut -c1-600
echo "== fused paths (1 GPU)";      timeout 400 python bench/fused_paths.py --reps 4 2>/dev/null | grep '^{' | tee $O/fused_1gpu.jsonl | cut -c1-300
echo "== fused paths, opt-in ones (1 GPU): GEMM -> reduce-scatter both landings, the batched lasso iteration"
timeout 400 python bench/fused_paths.py --reps 3 --experimental --skip xtr,hist,gram,transpose 2>/dev/null | grep '^{' | tee $O/fused_optin_1gpu.jsonl | cut -c1-300
echo "== stability bench";          timeout 300 python bench/stability_bench.py 2>&1 | tail -1 | tee $O/stability.json | cut -c1-600
echo "== pan-cancer bench";         timeout 600 python bench/pancancer_bench.py 2>&1 | tail -1 | tee $O/pancancer.json | cut -c1-600
echo "== gemm bench";               timeout 200 python bench/gemm_bench.py 2>/dev/null | tee $O/gemm_bench.jsonl | cut -c1-250
echo "== lasso phases";             timeout 200 python tools/prof_lasso_phases.py 2>&1 | tail -3 | tee $O/lasso_phases.jsonl | cut -c1-300
echo "== ncu: CTA-pair GEMM, histogram build, CTA-per-fit path kernel (after the plain runs above exited 0)"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gemm_tn_pair_kernel --launch-skip 1 --launch-count 1 \
  -o $O/prof_gemm_pair -f python tools/prof_kernels.py gemm 2>&1 | tail -2
timeout 400 ncu --set full --clock-control none --import-source on -k regex:hist_allreduce_kernel --launch-skip 1 --launch-count 1 \
  -o $O/prof_hist -f python tools/prof_kernels.py hist 2>&1 | tail -2
CML_LASSO_TEAM=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:lognet_path_cta_kernel --launch-count 1 \
  -o $O/prof_lasso_team -f python tools/time_lasso_reps.py 2>&1 | tail -2
