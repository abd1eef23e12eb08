from .validate import validate_path  # noqa: F401
from .clocks import ClockSampler  # noqa: F401
