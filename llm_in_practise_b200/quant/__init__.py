"""W4A16 post-training quantisation (GPTQ / AWQ / RTN), GPTQ + compressed-tensors checkpoint formats, front-ends."""
from .algorithms import awq_search_scale, capture, gptq_quantize
from .api import AWQModifier, GPTQModifier, QuantizationModifier, QuantizeConfig, tokenize_calibration
from .formats import (W4Weight, from_awq, from_compressed_tensors, from_gptq, pack_codes, quantize_rtn, to_compressed_tensors,
                      to_gptq)
from .frontends import GPTQModel, oneshot, perplexity, self_perplexity
from .pipeline import detect_quant_format, load_quantized, quantize_model, save_quantized
