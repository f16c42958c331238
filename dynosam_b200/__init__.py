"""dynosam_b200: B200-native backend Levenberg-Marquardt solve for DynoSAM graphs.

The product is libdynosam_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/dynosam_b200.h); this package is the thin host mirror of the reference's
backend interface for that path. There is no CPU fallback.
"""
from .api import (  # noqa: F401
    BackendParams, CameraPoseSymbol, DynamicLandmarkSymbol, DynoError, GpuLevenbergMarquardt, Graph, LmParams,
    ObjectMotionSymbol, ObjectPoseSymbol, PipelinedLevenbergMarquardt, Scene, StaticLandmarkSymbol, SyntheticConfig, ValuesKeyDoesNotExist,
    cantor_depair, cantor_pair, default_backend_params, default_lm_params, default_synthetic_config, key_decode, measure_fp64_tflops,
    sliding_window_schedule,
)
from . import graphfile  # noqa: E402,F401  (DynoSAM text graph file reader / writer)
from . import frontend_bson  # noqa: E402,F401  (the reference's recorded front-end packets, rgbd_frontend_output.bson)
from . import metrics  # noqa: E402,F401  (ATE / RPE / object-motion error against scene ground truth)
from . import postsolve  # noqa: E402,F401  (object pose propagation from the optimised motions, the reference logger CSV layout)
from . import sharded  # noqa: E402,F401  (sharded solve over the GPUs of one box: partition, rank shares, NCCL exchange inside the library)
from . import llworld  # noqa: E402,F401  (LLWorld formulation factors, batched evaluation on the device)
