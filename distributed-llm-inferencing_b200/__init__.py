"""B200-native sharded decoder forward for Distributed-LLM-Inferencing workers.

Host side in Python/PyTorch (device memory, streams, torch.distributed); all
arithmetic in hand-written sm_100a CUDA kernels behind the C-ABI of
include/dli.h.  Import as `dli_b200` (the directory name carries the
reference's hyphenated name and is not a valid module identifier).
"""
__version__ = "0.1.0"
