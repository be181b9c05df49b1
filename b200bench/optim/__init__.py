from .adamw import FlatAdamW, warmup_lr  # noqa: F401
