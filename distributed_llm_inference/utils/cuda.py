from distributed_llm_inference_b200.utils.cuda import make_inference_graphed_callable  # noqa: F401
