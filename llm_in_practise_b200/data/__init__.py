from .tokenizers import BPETokenizer, ByteTokenizer, CharTokenizer, load_or_train_bpe, load_pretrained_tokenizer  # noqa: F401
from .datasets import (DataCollatorForLanguageModeling, DataCollatorWithPadding, DistributedSampler,  # noqa: F401
                       SyntheticTokenDataset, TokenizedDataset, default_collate_pairs, prepare_text_data,
                       seed_worker, set_seed, synthetic_corpus)
from .chat import (CHATML_JINJA, build_messages, build_sft_features, convert_self_cognition_to_alpaca,  # noqa: F401
                   deepseek_prompt, load_self_cognition, preprocess_generic, render_chatml)
