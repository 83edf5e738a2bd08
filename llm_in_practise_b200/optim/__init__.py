from .adamw import (AdamW8bit, FusedAdamW, PagedAdamW8bit, clip_coef, clip_grad_norm_, grad_sumsq, scale_)  # noqa: F401
