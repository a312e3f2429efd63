from distributed_llm_inference_b200.utils.model import (convert_to_optimized_block, get_block_state_dict,  # noqa: F401
                                                          get_sharded_block_state_from_file, load_block)
