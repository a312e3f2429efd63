from distributed_llm_inference_b200.server.backend import InferenceBackend, TaskPool, BatchTensorDescriptor  # noqa: F401
