from .losses import ClipLoss, MSELoss  # noqa: F401
