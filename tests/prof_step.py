"""Profiling driver (not a test): two decode steps of the 7B-shaped span between cudaProfilerStart / Stop, for
`ncu --profile-from-start off`.   python tests/prof_step.py decode | decode_w8 | stage
  decode     32 layers, batch 16, context 512: LlamaBlock.graphed_decode (dli_span_decode_step)
  decode_w8  the same with row-wise int8 weights (gemm_w8_kernel)
  decode_a8  int8 weights and int8 activations (gemm_i8_kernel, LLM.int8)
  stage      8 layers as a one-GPU token-ring stage step (dli_stage_decode_step: embedding gather + flag wait at entry,
             final norm -> lm_head -> argmax -> flag release at exit)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from bench import build_span
from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.parallel.pipeline import PipelineStage
from distributed_llm_inference_b200.utils import synthetic as syn

mode = sys.argv[1] if len(sys.argv) > 1 else "decode"
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
shape = syn.SHAPES["7b"]
B, T = 16, 512
n_layers = 8 if mode == "stage" else 32
blk = build_span(shape, list(range(n_layers)), dev)
if mode == "decode_w8":
    blk.quantize_w8()
if mode == "decode_a8":
    blk.quantize_w8(activations="int8", threshold=5.0)
cache = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=2 * B * (T + 64) + 2048)
blk("g", syn.make_hidden(shape, B, T, seed=1, device=dev), past_key_value=cache)
x = syn.make_hidden(shape, B, 1, seed=2, device=dev)
if mode == "stage":
    cw = syn.make_decidable_client_weights(shape)
    stage = PipelineStage(blk, cache, 0, 1, dev, client={k: v.to(dev) for k, v in cw.items()})
    stage.enable_graphs(["g"], T + 32, batch=B, ring="tokens")
    tok = syn.make_prompt(shape, B, 1, seed=3, device=dev)[:, 0].to(torch.int32)
    step = lambda first: stage.run_round(["g"], [tok] if first else None, B, 1, 0, ring="tokens")
    step(True)
else:
    graph = blk.graphed_decode("g", cache, T + 32)
    step = lambda first: graph.step(x)
for _ in range(3):
    step(False)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(2):
    step(False)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled 2 steps of", mode)
