from .clocks import ClockSampler

__all__ = ["ClockSampler"]
