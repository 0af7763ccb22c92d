"""Micro-batching front for live traffic.

The reference router decides one request at a time inside
route_general_request (src/vllm_router/services/request_service/request.py:252-261):
many coroutines of one event loop each `await router.route_request(...)`.
The GPU kernels want batches.  MicroBatcher sits between the two: requests that
arrive within `max_delay_ms` (or until `max_batch` are waiting) are routed as one
batch, in arrival order, so the outcome equals the reference's one-by-one
routing of the same arrival sequence.  A batch needs one constant endpoint set,
so a request whose endpoint list differs from the pending ones first flushes
them (arrival order is kept across the split).

A flush does not run on the event loop: packing, the H2D copy, the kernels and the
wait for the decisions happen on ONE worker thread (batches stay in submission
order), and the waiters are resolved back on the loop when it is done -- the loop
keeps accepting requests (they form the next batch) while a batch is on the GPU.
"""
import asyncio
from concurrent.futures import ThreadPoolExecutor
from typing import Any, Callable, List, Optional, Sequence, Tuple


class MicroBatcher:
    def __init__(self, route_batch: Callable[[Sequence[Any], List[str]], List[str]],
                 max_batch: int = 4096, max_delay_ms: float = 1.0):
        """route_batch(endpoints, prompts) -> urls, e.g. PrefixAwareRouter.route_batch."""
        self._route_batch = route_batch
        self.max_batch = max_batch
        self.max_delay = max_delay_ms / 1e3
        self._pending: List[Tuple[str, asyncio.Future]] = []
        self._endpoints: Optional[Sequence[Any]] = None
        self._key: Optional[tuple] = None
        self._timer: Optional[asyncio.TimerHandle] = None
        self._worker = ThreadPoolExecutor(max_workers=1, thread_name_prefix="b200r-batch")
        self.batches_routed = 0
        self.requests_routed = 0

    @staticmethod
    def _endpoint_key(endpoints) -> tuple:
        return tuple(sorted(e.url if hasattr(e, "url") else e for e in endpoints))

    def _flush(self):
        if self._timer is not None:
            self._timer.cancel()
            self._timer = None
        if not self._pending:
            return
        pending, endpoints = self._pending, self._endpoints
        self._pending, self._endpoints, self._key = [], None, None
        loop = asyncio.get_running_loop()
        job = loop.run_in_executor(self._worker, self._route_batch, endpoints, [p for p, _ in pending])

        def done(fut):
            exc = fut.exception()
            if exc is not None:                           # every waiter sees the failure
                for _, w in pending:
                    if not w.done():
                        w.set_exception(exc)
                return
            self.batches_routed += 1
            self.requests_routed += len(pending)
            for (_, w), url in zip(pending, fut.result()):
                if not w.done():
                    w.set_result(url)
        job.add_done_callback(done)

    async def route(self, endpoints: Sequence[Any], prompt: str) -> str:
        loop = asyncio.get_running_loop()
        key = self._endpoint_key(endpoints)
        if self._pending and key != self._key:
            self._flush()                             # endpoint set changed: close the batch
        if not self._pending:
            self._endpoints, self._key = list(endpoints), key
            if self.max_delay > 0:
                self._timer = loop.call_later(self.max_delay, self._flush)
        fut = loop.create_future()
        self._pending.append((prompt, fut))
        if len(self._pending) >= self.max_batch or self.max_delay <= 0:
            self._flush()
        return await fut

    async def drain(self):
        self._flush()
