from distributed_llm_inference_b200.models.llama.cache import PartialLlamaSinkCache  # noqa: F401
