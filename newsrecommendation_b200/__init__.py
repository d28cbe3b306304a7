"""newsrecommendation_b200 — B200 (sm_100a) kernels behind the NRMS/NAML train-and-eval hot path of
patngnw/NewsRecommendation.

  _lib      ctypes binding of libnr_sm100.so (C ABI: include/nr_sm100.h)
  ops       torch.autograd.Function shims, one per reference operator
  model/    drop-in `model` plugin package (NRMS, NAML, model_utils)
  engine    training step (flat-bucket Adam + NCCL all-reduce) and the eval path
            (sharded news-vector cache, CSR impression scoring, ranking metrics)
  synthetic numpy generators of MIND-shaped inputs (no device code)

Every module that computes (`_lib`, `ops`, `model`, `engine`, `batcher`, `main`) loads the CUDA library on import and
fails loudly when it is missing: there is no fallback path.  The package object itself does not dlopen anything, so the
pure-numpy `synthetic` generators can be shared with the CPU reference arm of bench.py without mapping
libnr_sm100.so into that process."""
import importlib

__all__ = ['_lib', 'ops', 'engine', 'model', 'batcher', 'synthetic']


def __getattr__(name):
    if name in __all__:
        return importlib.import_module(f'{__name__}.{name}')
    raise AttributeError(f'module {__name__!r} has no attribute {name!r}')
