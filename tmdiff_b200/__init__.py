"""tmdiff_b200 -- B200-native sampling hot path of TMDiff (WavBEST denoiser + DPM-Solver++).

Python here is host glue over the C-ABI CUDA library (include/tmdiff_b200.h): it mirrors the
reference's module / sampler interfaces so the path drops into codgodtao/TMDiff unchanged.
"""
from . import _lib  # noqa: F401
from .networks import WavBEST, define_General  # noqa: F401
from .diffusion import GeneralDiffusion  # noqa: F401
from .dpm_solver import DPM_Solver, NoiseScheduleVP, model_wrapper  # noqa: F401
from . import ops  # noqa: F401
from . import operators  # noqa: F401
