"""llm_in_practise_b200 — a Blackwell (B200, sm_100a) native LLM training / finetuning /
int4-inference framework with the capabilities of the iKubernetes/llm-in-practise course code.

PyTorch is the tensor/autograd layer; the hot operators are hand-written CUDA kernels
(tcgen05 + TMEM + TMA) in ``csrc/`` exposed through a C ABI (``include/b200_ops.h``); multi-GPU
runs are one process per GPU over NCCL/NVLink with fused compute+collective kernels over
symmetric peer memory (``parallel/``).
"""
__version__ = "0.1.0"
