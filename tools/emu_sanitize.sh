#!/bin/bash
# Sanitizer run of the kernel sources on the host execution model (needs no GPU):
#   tools/emu_sanitize.sh            UBSan (aborts on the first undefined behaviour), then ASan + UBSan
# libcml_emu.so is rebuilt with the sanitizer flags (consensusml_b200/_build.py, CML_EMU_SANITIZE), the emulation tests run
# against it - every kernel, incl. the 2- / 3-rank fused paths and the tcgen05 / TMA GEMMs on the model of the asynchronous
# units - and the library is rebuilt plain.
# ASan sees the host buffers the "device" pointers refer to and the emulated dynamic shared memory (allocated at its exact
# size), so an out-of-bounds global or shared access of a kernel is reported with the kernel's source line.
set -e
cd "$(dirname "$0")/.."
mkdir -p profiles
OUT=profiles/sanitizer_emu.txt
echo "# tools/emu_sanitize.sh  $(date -u +%Y-%m-%dT%H:%MZ)  g++ $(g++ -dumpversion)" > $OUT
echo "== UBSan (-fsanitize=undefined -fno-sanitize-recover)" | tee -a $OUT
CML_EMU_SANITIZE=undefined python -m pytest tests/test_emu_kernels.py tests/test_emu_fused.py tests/test_emu_gemm.py tests/test_deseq.py -x -q 2>&1 | tail -2 | tee -a $OUT
echo "== ASan + UBSan (-fsanitize=address,undefined)" | tee -a $OUT
CML_EMU_SANITIZE=address LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0:detect_stack_use_after_return=0:verify_asan_link_order=0 \
  python -m pytest tests/test_emu_kernels.py tests/test_emu_fused.py tests/test_emu_gemm.py tests/test_deseq.py -x -q 2>&1 | tail -2 | tee -a $OUT
python -c "from consensusml_b200 import _build; _build.build_emu()"      # back to the plain build
