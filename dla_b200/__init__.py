"""dla_b200 — a Blackwell (B200 / sm_100a) native LLM alignment framework.

Phases: SFT -> reward model -> DPO / PPO-RLHF -> distillation -> eval, driven by
``config/*.yaml`` and ``scripts/launch_*.sh`` (same surface as the reference,
``R/README.md:40-46``).  The Python layer is PyTorch; the hot ops are hand-written
sm_100a CUDA kernels in ``csrc/`` exposed through a plain C ABI
(``include/dla_b200.h``) and loaded with ctypes by :mod:`dla_b200.ops.native`.
"""

__version__ = "0.1.0"
