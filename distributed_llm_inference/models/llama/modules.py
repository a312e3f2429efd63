from distributed_llm_inference_b200.models.llama.modules import OptimizedLlamaDecoderLayer, OptimizedLlamaInferenceAttention  # noqa: F401
