"""B200-native span forward for distributed-llm-inference (host API mirrors the reference's)."""
__version__ = "0.1.0"
