from .lora import (ALL_LINEAR, LoraConfig, PeftModel, apply_lora, get_peft_model,  # noqa: F401
                   prepare_model_for_kbit_training)
