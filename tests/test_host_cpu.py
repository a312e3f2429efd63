"""CPU-only checks: the C-ABI library loads and exports every symbol include/dli_b200.h declares (no compute
calls without a GPU), argument validation that needs no device, and the host logic of the paged cache against
the oracle's restatement of the reference cache."""
import ctypes as C
import os
import re

import pytest
import torch

from distributed_llm_inference_b200.models.llama import LlamaBlock, PartialLlamaSinkCache
from distributed_llm_inference_b200.server.backend import BatchTensorDescriptor, InferenceBackend
from distributed_llm_inference_b200.utils import synthetic as syn
from oracle import span_oracle as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from distributed_llm_inference_b200 import _ffi

    header = open(os.path.join(ROOT, "include", "dli_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(dli_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(_ffi.EXPORTS), declared ^ set(_ffi.EXPORTS)
    for name in declared:
        assert hasattr(_ffi.lib, name), f"{name} not exported by {_ffi.LIB_PATH}"
    assert _ffi.lib.dli_version() == 1


def test_host_side_argument_validation_needs_no_gpu():
    from distributed_llm_inference_b200 import _ffi

    out = C.c_void_p()
    bad = _ffi.SpanDesc(768, 2048, 12, 5, 64, 2, 0, 1e-6, 1e-6, 2.0)          # 12 heads, 5 kv heads
    assert _ffi.lib.dli_span_create(C.byref(bad), C.byref(out)) == -1
    assert b"n_heads must be a multiple" in _ffi.lib.dli_last_error()
    bad = _ffi.SpanDesc(768, 2048, 12, 12, 80, 2, 0, 1e-6, 1e-6, 2.0)         # head_dim 80
    assert _ffi.lib.dli_span_create(C.byref(bad), C.byref(out)) == -1
    ok = _ffi.SpanDesc(768, 2048, 12, 12, 64, 2, 0, 1e-6, 1e-6, 2.0)
    assert _ffi.lib.dli_span_create(C.byref(ok), C.byref(out)) == 0
    assert _ffi.lib.dli_span_workspace_bytes(out, 16, 16, 1024) > 0
    assert _ffi.lib.dli_span_workspace_bytes(out, 0, 16, 1024) == 0
    w = _ffi.LayerWeights(None, None, None, None, None, None)
    assert _ffi.lib.dli_span_bind_weights(out, 0, C.byref(w)) == -1
    assert _ffi.lib.dli_span_bind_weights(out, 7, C.byref(w)) == -1
    assert _ffi.lib.dli_span_destroy(out) == 0
    with pytest.raises(_ffi.DliError):
        _ffi.check(-3)


def test_block_keeps_reference_parameter_names_and_refuses_cpu():
    shp = syn.SHAPES["test_gqa"]
    blk = LlamaBlock(shp.to_hf_config(), [4, 5])
    names = set(blk.state_dict().keys())
    want = {f"layers.{j}.{n}.weight" for j in range(2) for n in syn.LINEARS + syn.NORMS}
    assert names == want
    W = syn.make_span_weights(shp, [4, 5])
    blk.load_state_dict({f"layers.{j}.{n}.weight": t for j, w in enumerate(W) for n, t in w.items()})
    for layer in blk.layers:
        layer.fuse()
        a = layer.self_attn
        assert torch.equal(layer._wqkv, torch.cat([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight]))
        assert a.k_proj.weight.data_ptr() == layer._wqkv.data_ptr() + a.q_proj.weight.numel() * 2   # views, not copies
        assert torch.equal(layer._wgu, torch.cat([layer.mlp.gate_proj.weight, layer.mlp.up_proj.weight]))
    assert torch.equal(blk.layers[1].self_attn.q_proj.weight, W[1]["self_attn.q_proj"])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        blk("g", torch.zeros(1, 3, shp.hidden_size, dtype=torch.bfloat16))


def test_rope_table_matches_oracle_tables():
    from distributed_llm_inference_b200.models.llama.model import rope_table

    pos = torch.arange(0, 300)[None]
    for mode in ("literal", "float"):
        cos, sin = so.rope_tables(pos, 64, 10000.0, mode, torch.float32)
        tab = rope_table(64, 10000.0, 300, mode, "cpu")
        assert torch.equal(tab[0], cos[0].float()) and torch.equal(tab[1], sin[0].float())


def test_paged_cache_update_matches_oracle_cache():
    """The tensor-level ``update`` API (cache.py:64-135) on CPU tensors against the oracle cache: two
    generations with different batch sizes, appends that cross page boundaries."""
    n_layers, KV, hd = 2, 2, 8
    c = PartialLlamaSinkCache(4096, 4, page_tokens=4, capacity_tokens=256)
    c.bind(n_layers, KV, hd, "cpu", dtype=torch.float32)
    o = so.SinkCacheOracle(4096, 4)
    g = torch.Generator().manual_seed(0)
    for gid, B, n in [("a", 1, 3), ("b", 3, 6), ("a", 1, 1), ("a", 1, 7), ("b", 3, 1), ("b", 3, 1)]:
        for layer in range(n_layers):
            k, v = torch.randn(B, KV, n, hd, generator=g), torch.randn(B, KV, n, hd, generator=g)
            K, V = c.update(k, v, layer, {"generation_id": gid})
            Ko, Vo = o.update(k, v, layer, {"generation_id": gid})
            assert torch.equal(K, Ko) and torch.equal(V, Vo)
        assert c.get_seq_length(0, gid) == o.get_seq_length(0, gid)
    assert c.get_seq_length(5, "a") == 0 and c.get_seq_length(0, "nobody") == 0
    used = c.n_pages - c.free_pages
    assert used == 3 + 3 * 2                                # a: 11 tokens -> 3 pages; b: 8 tokens x 3 seqs -> 6
    c.close("b")
    assert c.n_pages - c.free_pages == 3
    with pytest.raises(ValueError):
        c.get_seq_length(0, None)                           # cache.py:51-52
    with pytest.raises(ValueError):
        c.update(k, v, 0, {})                               # cache.py:71-72
    with pytest.raises(ValueError):
        c.open("a", 2)                                      # batch mismatch for a live generation


def test_paged_cache_limits_are_loud():
    c = PartialLlamaSinkCache(10, 2, page_tokens=4, capacity_tokens=16)
    c.bind(1, 1, 8, "cpu", dtype=torch.float32)
    c.reserve("g", 1, 8)
    c.advance("g", 8)
    assert c.needs_eviction("g", 2) and not c.needs_eviction("g", 1)    # cache.py:107: append while n + len < window
    with pytest.raises(ValueError, match="evict first"):
        c.reserve("g", 1, 11)
    with pytest.raises(ValueError, match="cos/sin history"):      # update() evicts, and eviction re-rotates keys
        c.update(torch.zeros(1, 1, 2, 8), torch.zeros(1, 1, 2, 8), 0, {"generation_id": "g"})
    c2 = PartialLlamaSinkCache(4096, 2, page_tokens=4, capacity_tokens=8)
    c2.bind(1, 1, 8, "cpu", dtype=torch.float32)
    with pytest.raises(RuntimeError, match="exhausted"):
        c2.reserve("g", 3, 8)


def test_inference_backend_contract():
    """backend.py:12-51: schemas, flat-tuple forward under no_grad, single pool, backward refused."""

    class Echo(torch.nn.Module):
        def forward(self, generation_id, hidden, cache_position=None, past_key_value=None):
            return (hidden * 2, )

    sch = BatchTensorDescriptor((5, 8), torch.float32)
    be = InferenceBackend("blk", Echo(), args_schema=(sch,))
    assert be.outputs_schema == (sch,) and be.get_pools() is be.inference_pool
    x = torch.ones(3, 5, 8, requires_grad=True)
    (y,) = be.session("gen").forward(x)
    assert torch.equal(y, x * 2) and not y.requires_grad
    assert be.inference_pool.submit_task(x, generation_id="g2")[0].shape == x.shape
    with pytest.raises(NotImplementedError):
        be.backward(x)
    with pytest.raises(NotImplementedError):
        be.on_backward(3)
    with pytest.raises(AssertionError):
        InferenceBackend("x", Echo())


def test_load_block_from_local_sharded_checkpoint(tmp_path):
    """utils/model.py:16-90 contract, offline: only the shards holding the span's layers are read, parameters keep
    the reference's names, floating tensors are cast to the span dtype."""
    import json

    import safetensors.torch

    from distributed_llm_inference_b200.utils.model import get_block_state_dict, load_block

    shp = syn.SHAPES["test_gqa"]
    shp.to_hf_config().save_pretrained(tmp_path)
    weight_map, shards = {}, {0: {}, 1: {}}
    per_layer = {}
    for idx in range(shp.num_hidden_layers):
        w = syn.make_layer_weights(shp, idx, dtype=torch.float16, norm_jitter=0.1)     # checkpoints are often fp16
        per_layer[idx] = w
        for name, t in w.items():
            key = f"model.layers.{idx}.{name}.weight"
            shard = idx % 2 if "mlp" in name else (idx + 1) % 2                       # a layer spans two shards
            shards[shard][key] = t
            weight_map[key] = f"model-{shard:05d}-of-00002.safetensors"
    shards[0]["model.embed_tokens.weight"] = torch.zeros(4, 4)
    weight_map["model.embed_tokens.weight"] = "model-00000-of-00002.safetensors"
    for s, tensors in shards.items():
        safetensors.torch.save_file(tensors, str(tmp_path / f"model-{s:05d}-of-00002.safetensors"))
    (tmp_path / "model.safetensors.index.json").write_text(json.dumps({"metadata": {}, "weight_map": weight_map}))

    state = get_block_state_dict(str(tmp_path), 2)
    assert set(state) == {f"{n}.weight" for n in syn.LINEARS + syn.NORMS}
    blk = load_block(str(tmp_path), [1, 3], rope="float", attn_implementation="eager")
    assert [layer.layer_idx for layer in blk.layers] == [1, 3]
    for j, idx in enumerate([1, 3]):
        for name, t in per_layer[idx].items():
            got = blk.state_dict()[f"layers.{j}.{name}.weight"]
            assert got.dtype == torch.bfloat16 and torch.equal(got, t.to(torch.bfloat16))
    with pytest.raises(FileNotFoundError):
        get_block_state_dict(str(tmp_path / "nowhere"), 0)
    with pytest.raises(NotImplementedError):
        load_block(str(tmp_path), [0], use_quantized=True)
