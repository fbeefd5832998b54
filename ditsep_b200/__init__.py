"""ditsep_b200: the DiffSep predictor-corrector separation sampler on B200 (sm_100a).

Mirrors the reference's import paths for this path under one package:
``ditsep_b200.sdes`` (src/sdes), ``ditsep_b200.models.diffsep`` (src/models/diffsep),
``ditsep_b200.diffsep`` (src/diffsep.py), ``ditsep_b200.inference.separate``
(src/inference/separate.py), ``ditsep_b200.utils`` (src/utils).  All numerics run in the
hand-written CUDA library ``libditsep_b200.so`` (C ABI in include/ditsep_b200.h).
"""
__version__ = "0.1.0"
