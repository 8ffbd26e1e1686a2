"""Native runtime pieces that sit between the host program and the captured step graphs."""
from .feeder import HostFeeder, Ticket  # noqa: F401
