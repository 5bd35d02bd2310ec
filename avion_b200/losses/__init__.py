from .losses import ClipLoss, MaxMarginRankingLoss, MSELoss  # noqa: F401
