"""Serving: continuous-batching engine, OpenAI-compatible server, model router, cache gateway, moderation wrapper, load generator."""
from .engine import EngineArgs, LLMEngine, Request
