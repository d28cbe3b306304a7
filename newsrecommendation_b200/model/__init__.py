"""Drop-in replacement for the reference's `src/model` plugin package
(importlib.import_module(f'model.{args.model}'), src/main.py:63,169).
Modules: NRMS, NAML, model_utils — same class names, constructors, forward
signatures and state_dict keys; all arithmetic runs in libnr_sm100.so."""
