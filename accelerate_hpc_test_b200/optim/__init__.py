from .adamw import FusedAdamW, grad_global_norm
