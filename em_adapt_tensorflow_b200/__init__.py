"""em_adapt_tensorflow_b200 -- B200-native (sm_100a) EM-Adapt E-step.

Only the hot path of xtudbxk/em-adapt-tensorflow lives here: the latent-label E-step
(estep/weaklabel.cpp behind estep.py) and the softmax-CE that consumes its labels
(deeplab.py:109-124, 180-186).  The product is csrc/ -> lib/libweaklabel_cuda.so (a C-ABI
library, include/weaklabel_cuda.h); the Python modules are thin host-side mirrors of the
reference's own interface for this path.
"""
from ._lib import WeakLabelError, lib_path, load            # noqa: F401
from .estep import estep, estep_c                            # noqa: F401
from .plan import EStepPlan, OrderStream, draws_for, label_set, label_set_device, softmax_ce, fixed_bias_labels, upsample_bilinear   # noqa: F401
from .shard import shard_range, draws_before, positioned_stream, draw_order   # noqa: F401
