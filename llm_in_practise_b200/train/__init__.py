"""Hand-written training loops (pretraining runner) behind the recipes."""
from .pretrain import PretrainConfig, run_pretraining
