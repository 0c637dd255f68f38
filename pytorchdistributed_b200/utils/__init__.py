from .clocks import ClockSampler
from .metrics import JsonlWriter, StepMeter
from .nvtx import nvtx_range

__all__ = ["ClockSampler", "JsonlWriter", "StepMeter", "nvtx_range"]
