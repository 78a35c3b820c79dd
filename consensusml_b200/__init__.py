"""consensusml_b200 - Blackwell-native consensus biomarker selection.

Layer map: ``csrc/`` sm_100a kernels behind a C ABI -> ``ops`` -> ``stats`` +
``models`` -> ``consensus`` -> ``io`` / ``api`` / ``cli``; ``parallel`` holds the NCCL bootstrap and the
fused compute+collective paths over NVLink peer memory.
"""
__version__ = "0.1.0"
