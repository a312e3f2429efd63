from distributed_llm_inference_b200.models.llama import *  # noqa: F401,F403
