from .losses import ClipLoss  # noqa: F401
