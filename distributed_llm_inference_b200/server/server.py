"""``Server`` — the loop that hosts one span of a model and serves requests for it.

Reference: distributed_llm_inference/server/server.py:4-24 is commented pseudo-code (choose block ids, build the
module, then ``runtime.wait()`` in a loop until the module is unhealthy or the swarm should be rebalanced, finally
restart), shadowing the hivemind ``Server`` it imports at :1.  This version is that loop made real for one box:

  * ``run()`` builds the ``InferenceWorker`` for ``[block_index_start, block_index_end)`` (server/worker.py), then
    takes tasks from a thread-safe queue and runs them one at a time on the worker's backend pool -- calls into a span
    are serialised, as hivemind's Runtime thread serialises them;
  * every task names its session (``generation_id``) out of band of the tensors;
  * after every task it checks ``is_healthy()`` (a failed kernel launch or a poisoned CUDA context) and
    ``should_rebalance()`` (always False on one box; a swarm scheduler would override it) and, as in the pseudo-code,
    leaves the inner loop to rebuild the worker (``restart``) when either says so;
  * ``submit()`` is the client side of the queue and returns a ``concurrent.futures.Future``; ``shutdown()`` ends
    ``run()``.
Transport between machines (hivemind in the reference's imports) is outside this repository: on one NVSwitch box the
stages are connected by ``parallel/pipeline.py`` instead.
"""
from __future__ import annotations

import queue
import threading
from concurrent.futures import Future
from typing import Callable, Optional

import torch

_STOP = object()


class Server:
    def __init__(self, model: str, block_index_start: int, block_index_end: int, *,
                 worker_factory: Optional[Callable] = None, **worker_kwargs):
        """``worker_factory(model, start, end, **kw)`` builds the object that hosts the span (default:
        ``InferenceWorker``); tests pass a fake."""
        self.model, self.block_index_start, self.block_index_end = model, block_index_start, block_index_end
        self.worker_kwargs = worker_kwargs
        if worker_factory is None:
            from distributed_llm_inference_b200.server.worker import InferenceWorker as worker_factory
        self.worker_factory = worker_factory
        self.worker = None
        self.tasks: "queue.Queue" = queue.Queue()
        self.ready = threading.Event()
        self.restarts = 0
        self.served = 0
        self._failed = False
        self._stopping = False

    # ---- hooks named by the reference's pseudo-code (server.py:18-19) -------------------------------------
    def is_healthy(self) -> bool:
        return not self._failed

    def should_rebalance(self) -> bool:
        return False

    # ---- client side ---------------------------------------------------------------------------------------
    def submit(self, generation_id: str, hidden_states: torch.Tensor,
               cache_position: Optional[torch.Tensor] = None) -> Future:
        fut: Future = Future()
        self.tasks.put((generation_id, hidden_states, cache_position, fut))
        return fut

    def shutdown(self) -> None:
        self._stopping = True
        self.tasks.put(_STOP)

    # ---- the loop (server.py:5-24) ---------------------------------------------------------------------------
    def run(self) -> None:
        while not self._stopping:
            # "self.module = new Module(...)": load the span for this server's block range
            self.worker = self.worker_factory(self.model, self.block_index_start, self.block_index_end, **self.worker_kwargs)
            self._failed = False
            self.ready.set()
            try:
                while True:
                    task = self.tasks.get()                       # "self.module.runtime.wait()"
                    if task is _STOP:
                        return
                    gid, x, cp, fut = task
                    if not fut.set_running_or_notify_cancel():
                        continue
                    try:
                        pool = self.worker.run()                  # the backend's inference pool
                        out = pool.submit_task(x, generation_id=gid, cache_position=cp)[0]
                        fut.set_result(out)
                        self.served += 1
                    except Exception as e:  # noqa: BLE001 - the client gets the error, the loop decides what to do next
                        fut.set_exception(e)
                        self._failed = self._is_fatal(e)
                    if not self.is_healthy() or self.should_rebalance():
                        break
            finally:
                self.ready.clear()
                if not self._stopping:                            # "self.module.restart()"
                    self.restarts += 1
                    self.worker = None

    @staticmethod
    def _is_fatal(e: Exception) -> bool:
        """Errors that leave the span unusable: a failed CUDA call inside the library (its workspace invariants are
        gone) or a CUDA error raised by torch.  Argument errors (wrong dtype, missing generation id) are the caller's."""
        from distributed_llm_inference_b200._ffi import DliError

        if isinstance(e, DliError):
            return "error -2" in str(e)
        return isinstance(e, RuntimeError) and "CUDA" in str(e)
