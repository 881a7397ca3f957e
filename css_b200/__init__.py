"""css_b200: B200-native CSS hot path behind the desh2608/css Python surface.

Sub-packages mirror the reference: ``css_b200.models`` (css/models), ``css_b200.executor``
(css/executor), ``css_b200.objectives`` (css/objectives), ``css_b200.utils.model_util``
(css/utils/model_util.py).  Everything numeric runs in libcss_b200.so (hand-written sm_100a
kernels, include/css_b200.h); there is no CPU or PyTorch fallback.
"""
__version__ = "0.1.0"
