
from .accuracy import StreamingAccuracy, accuracy  # noqa: F401
