from distributed_llm_inference_b200.models.llama.model import LlamaBlock  # noqa: F401
