"""Import-path alias: ``distributed_llm_inference.*`` resolves to the B200-native implementation in
``distributed_llm_inference_b200`` so that code written against the reference's module paths
(models/llama/{model,modules,cache}.py, utils/cuda.py, server/backend.py) runs unchanged."""
