from distributed_llm_inference_b200.server.worker import Block, InferenceWorker  # noqa: F401
