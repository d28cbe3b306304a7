"""newsrecommendation_b200 — B200 (sm_100a) kernels behind the NRMS/NAML train-and-eval
hot path of patngnw/NewsRecommendation.

  _lib      ctypes binding of libnr_sm100.so (C ABI: include/nr_sm100.h)
  ops       torch.autograd.Function shims, one per reference operator
  model/    drop-in `model` plugin package (NRMS, NAML, model_utils)
  engine    training step (flat-bucket Adam + NCCL all-reduce) and the eval path
            (sharded news-vector cache, CSR impression scoring, ranking metrics)

Importing the package loads the CUDA library and fails loudly when it is missing.
"""
from . import _lib  # noqa: F401  (raises ImportError when libnr_sm100.so is absent)

__all__ = ['_lib']
