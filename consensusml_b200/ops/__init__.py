"""Kernel-level operations.  Every op dispatches on the tensor's device: CUDA tensors go to the
hand-written sm_100a kernels in ``libcml_b200.so``; CPU tensors go to the float64 oracle
(``libcml_host.so`` / NumPy), which is what the non-GPU test-suite exercises."""
