"""The round-2 kernels on a real GPU: the same backend-agnostic cases that the CPU suite runs on the host execution model
(tests/kernel_cases.py), here with CUDA tensors and the sm_100a library.  The file name sorts after every other test file
except the allocation-free entry point's, so a stop-at-first-failure run has already covered the round-1 kernels."""
import pytest

import kernel_cases as KC

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", [n for n in dir(KC) if n.startswith("case_")])
def test_round2_kernel_on_cuda(case):
    getattr(KC, case)(KC.CudaBackend())
