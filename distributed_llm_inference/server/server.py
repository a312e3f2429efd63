from distributed_llm_inference_b200.server.server import Server  # noqa: F401
