"""Serving: continuous-batching engine, OpenAI-compatible server, moderation wrapper, load generator."""
from .engine import EngineArgs, LLMEngine, Request
