"""mtldas_b200 -- B200-native (sm_100a) implementation of the MTL-DAS hot path.

Host side of the C-ABI library ``libmtldas_b200.so`` (include/mtldas.h): the drop-in
``MTL_Net`` / ``Single_Task_Net`` modules keep the reference's constructors, forward
signature and state_dict layout (/root/reference/model/modelA_MTL.py:53-174,
modelB_singleTask.py:53-178) while their compute runs in hand-written CUDA kernels.
There is no CPU fallback: without the built library and a B200 the compute calls raise.
"""
from ._lib import lib, MtldasError, library_path  # noqa: F401
from .engine import (EvalPlan, TrainPlan, FlatAdam, EvalMetrics, conv2d_nchw, conv2d_nchw_static, nll_loss,  # noqa: F401
                     umma_selftest)
from .nets import MTL_Net, Single_Task_Net  # noqa: F401
from . import synth  # noqa: F401
from .prefetch import DevicePrefetcher  # noqa: F401

__all__ = ["MTL_Net", "Single_Task_Net", "EvalPlan", "TrainPlan", "FlatAdam", "EvalMetrics", "conv2d_nchw", "umma_selftest",
           "lib", "MtldasError", "DevicePrefetcher"]
