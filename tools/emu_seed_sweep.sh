#!/bin/bash
# Race / uninitialised-read sweep on the host execution model: the emulation tests again under shifted seeds (thread order,
# shared-memory garbage, completion order of TMA / tcgen05 operations) and under both fixed poison patterns.
#   tools/emu_seed_sweep.sh [shift ...]        default shifts: 101 202 303
set -e
cd "$(dirname "$0")/.."
OUT=profiles/emu_seed_sweep.txt
echo "# tools/emu_seed_sweep.sh  $(date -u +%Y-%m-%dT%H:%MZ)" > $OUT
FILES="tests/test_emu_kernels.py tests/test_emu_fused.py tests/test_emu_gemm.py tests/test_deseq.py"
for s in ${@:-101 202 303}; do
  echo "== CML_EMU_SEED_SHIFT=$s" | tee -a $OUT
  CML_EMU_SEED_SHIFT=$s python -m pytest $FILES -x -q -p no:cacheprovider 2>&1 | tail -1 | tee -a $OUT
done
echo "== CML_EMU_POISON_EMPTY=1 (torch.empty returns NaN / large integers, as device memory is not zero)" | tee -a $OUT
CML_EMU_POISON_EMPTY=1 python -m pytest $FILES tests/test_emu_dist.py -x -q -p no:cacheprovider 2>&1 | tail -1 | tee -a $OUT
for p in ff 00; do
  echo "== CML_EMU_POISON=$p CML_EMU_SEED_SHIFT=17" | tee -a $OUT
  CML_EMU_POISON=$p CML_EMU_SEED_SHIFT=17 python -m pytest $FILES -x -q -p no:cacheprovider 2>&1 | tail -1 | tee -a $OUT
done
