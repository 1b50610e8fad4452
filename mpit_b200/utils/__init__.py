"""Utilities: loggers, confusion matrix, timers (the reference's optim.Logger / optim.ConfusionMatrix /
sys.clock buckets: /root/reference/asyncsgd/goot.lua:24-26,93-97,236-244)."""
from .logging import Logger, ConfusionMatrix, Timers, progress
