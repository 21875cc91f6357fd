"""Engine adapter that puts the paged decode path behind the reference's engine surface."""
from .paged_engine import ByteTokenizer, PagedLLMEngine

ENGINE_REGISTRY = {"llm_paged_b200": PagedLLMEngine}

__all__ = ["PagedLLMEngine", "ByteTokenizer", "ENGINE_REGISTRY"]
