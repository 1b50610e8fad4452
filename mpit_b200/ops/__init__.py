"""Python front-ends of the sm_100a kernels (thin ctypes wrappers over include/mpit_b200.h).

Every wrapper takes CUDA tensors, launches on the current stream and raises if the native
library is missing -- there is no eager PyTorch fallback on this path.
"""
from .optim_ops import (RULES, local_step, rule_step, downpour_accum, grad_reg_clip, map_copy,
                        rule_hyper)
from .nn_ops import (ConvGeom, conv_fwd_ref, conv_dgrad_ref, conv_wgrad_ref, act_bwd, colsum,
                     prep_u8, nll_fwd_bwd, classifier_head, classifier_head_scratch)
