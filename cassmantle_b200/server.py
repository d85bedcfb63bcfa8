"""MicroBatcher -- the request micro-batcher north_star places in src/server.py.

The reference scores each request inside its own coroutine: ``compute_scores`` is ``async``
but never awaits (src/backend.py:312-317), so every ``POST /compute_score`` (main.py:113-120)
blocks the event loop for its two words and nothing is ever batched.  This class coalesces
the ``compute_scores`` calls that are pending at the same time into ONE device batch:

    scores = await self._batcher.compute_scores(pairs)        # replaces src/server.py:73

* a request joins the open batch; the batch is flushed when it holds ``max_batch`` pairs or
  ``max_delay_s`` after its first request arrived, whichever comes first;
* the flush runs ``scorer.score_pairs`` on a worker thread, so the loop keeps accepting
  requests (they form the next batch) while the GPU works;
* every caller is resolved with only its own keys, in its own order, as ``str(score)``
  exactly like src/backend.py:316; a scorer failure is raised in every caller of that batch.

``scorer`` is anything with ``score_pairs(list[(guess, answer)]) -> list`` --
cassmantle_b200.backend.Backend in production.
"""
from __future__ import annotations

import asyncio
from concurrent.futures import ThreadPoolExecutor
from typing import Dict, List, Optional, Tuple


class MicroBatcher:
    def __init__(self, scorer, max_batch: int = 4096, max_delay_s: float = 0.002) -> None:
        if max_batch < 1:
            raise ValueError("max_batch must be positive")
        self.scorer = scorer
        self.max_batch = max_batch
        self.max_delay_s = max_delay_s
        self._pending: List[Tuple[List[str], List[Tuple[str, str]], asyncio.Future]] = []
        self._pending_pairs = 0
        self._timer: Optional[asyncio.TimerHandle] = None
        self._executor = ThreadPoolExecutor(max_workers=1, thread_name_prefix="cm-score")
        self._tasks = set()
        self.batches = 0            # device batches launched (observability / tests)
        self.requests = 0

    async def compute_scores(self, data: Dict[str, Dict[str, str]]) -> Dict[str, str]:
        keys = list(data.keys())
        # a malformed payload fails its own request, as in the reference, never the batch
        pairs = [(data[k]["input"], data[k]["answer"]) for k in keys]
        if not keys:
            return {}
        loop = asyncio.get_running_loop()
        fut: asyncio.Future = loop.create_future()
        self._pending.append((keys, pairs, fut))
        self._pending_pairs += len(pairs)
        self.requests += 1
        if self._pending_pairs >= self.max_batch:
            self._flush()
        elif self._timer is None:
            self._timer = loop.call_later(self.max_delay_s, self._flush)
        return await fut

    def _flush(self) -> None:
        if self._timer is not None:
            self._timer.cancel()
            self._timer = None
        if not self._pending:
            return
        batch, self._pending, self._pending_pairs = self._pending, [], 0
        task = asyncio.get_running_loop().create_task(self._run(batch))
        self._tasks.add(task)
        task.add_done_callback(self._tasks.discard)

    async def _run(self, batch) -> None:
        flat = [p for _, pairs, _ in batch for p in pairs]
        self.batches += 1
        try:
            vals = await asyncio.get_running_loop().run_in_executor(
                self._executor, self.scorer.score_pairs, flat)
        except Exception as exc:                       # noqa: BLE001 - forwarded to the callers
            for _, _, fut in batch:
                if not fut.done():
                    fut.set_exception(exc)
            return
        pos = 0
        for keys, pairs, fut in batch:
            out: Dict[str, str] = {}
            for k, v in zip(keys, vals[pos:pos + len(pairs)]):
                out.update({k: str(v)})                # src/backend.py:316
            pos += len(pairs)
            if not fut.done():
                fut.set_result(out)

    async def drain(self) -> None:
        """Flush what is pending and wait for every batch in flight."""
        self._flush()
        if self._tasks:
            await asyncio.gather(*list(self._tasks), return_exceptions=True)

    def close(self) -> None:
        self._executor.shutdown(wait=True)
