"""Server loop (server/server.py) on CPU with a fake worker: tasks are served in order with their generation id out of
band, client errors reach the caller without killing the loop, a fatal error rebuilds the worker (the reference's
``restart``, server.py:23), ``shutdown`` ends ``run``."""
import threading

import pytest
import torch

from distributed_llm_inference_b200.server.backend import BatchTensorDescriptor, InferenceBackend
from distributed_llm_inference_b200.server.server import Server


class _Echo(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.calls = []

    def forward(self, generation_id, hidden, cache_position=None, past_key_value=None):
        self.calls.append((generation_id, None if cache_position is None else int(cache_position[0])))
        if generation_id == "bad":
            raise ValueError("generation_id not provided")
        if generation_id == "fatal":
            raise RuntimeError("CUDA error: an illegal memory access was encountered")
        return (hidden * 2,)


class _FakeWorker:
    built = 0

    def __init__(self, model, start, end, **kw):
        type(self).built += 1
        self.module = _Echo()
        self.backend = InferenceBackend(f"span.{start}:{end}", self.module, args_schema=(BatchTensorDescriptor((1, 4), torch.float32),))

    def run(self):
        return self.backend.get_pools()


def test_server_loop_serves_restarts_and_stops():
    _FakeWorker.built = 0
    srv = Server("unused", 2, 5, worker_factory=_FakeWorker)
    t = threading.Thread(target=srv.run, daemon=True)
    t.start()
    assert srv.ready.wait(10)
    x = torch.ones(2, 1, 4)
    a = srv.submit("g1", x, torch.tensor([7]))
    b = srv.submit("g2", x * 3)
    assert torch.equal(a.result(10), x * 2) and torch.equal(b.result(10), x * 6)
    first = srv.worker
    assert first.module.calls == [("g1", 7), ("g2", None)]
    with pytest.raises(ValueError):
        srv.submit("bad", x).result(10)               # the caller's error: the loop goes on with the same worker
    assert torch.equal(srv.submit("g1", x).result(10), x * 2) and srv.worker is first and srv.restarts == 0
    with pytest.raises(RuntimeError):
        srv.submit("fatal", x).result(10)             # the span is unusable: restart
    assert torch.equal(srv.submit("g3", x).result(10), x * 2)
    assert srv.restarts == 1 and _FakeWorker.built == 2 and srv.worker is not first
    srv.shutdown()
    t.join(10)
    assert not t.is_alive() and srv.served == 4
