
from .trace import StepTimer, device_time_ms, nvtx_range  # noqa: F401
