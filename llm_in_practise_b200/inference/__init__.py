"""Generation: KV-cache runner, sampling, streaming, chat."""
from .generate import GenerationConfig, Generator, KVCache, TextIteratorStreamer, chat, sample_next, stream_generate
