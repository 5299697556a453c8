"""Host-side frame I/O around the online loop: the reference reads, converts, writes and scores every
frame synchronously between two GPU phases (blind_denoising.py:170-203 reads, :233-253 writes and PSNR).  At ~5
frames/s a 1080p PNG decode (~25 ms) plus encode (~90 ms) plus the reference-frame decode for the PSNR would cost a
quarter of the frame time, so both sides move to host threads:

* ``Prefetcher``  -- ``seq[k]`` hands back item k and keeps the next ``depth`` reads running on a small pool;
* ``OrderedSink`` -- jobs run on ONE worker in submission order (the order of ``plot_psnr.txt`` lines, of the
  ``print(i)`` / ``print(psnr)`` pairs and of the output files is the reference's), at most ``depth`` in flight.

Results are the reference's byte for byte: the same functions run, only on another thread.  Errors surface on the
calling thread (at ``seq[k]`` / at the next ``submit`` or ``close``) -- nothing is swallowed.
"""
import threading
from collections import deque
from concurrent.futures import ThreadPoolExecutor


class Prefetcher:
    """A read-ahead view of ``read(k)`` for k in [0, n): sequential consumers (OnlineF2F.run_sequence reads frame
    indices in increasing order, `lookahead` at a time) find their item already decoded."""

    def __init__(self, read, n, depth=4, workers=2):
        self._read, self._n, self._depth = read, int(n), max(0, int(depth))
        self._pool = ThreadPoolExecutor(max_workers=max(1, int(workers)), thread_name_prefix="f2f-read")
        self._pending = {}
        self._lock = threading.Lock()

    def __len__(self):
        return self._n

    def _schedule(self, k):
        if 0 <= k < self._n and k not in self._pending:
            self._pending[k] = self._pool.submit(self._read, k)

    def __getitem__(self, k):
        if k < 0:
            k += self._n
        if not 0 <= k < self._n:
            raise IndexError(k)
        with self._lock:
            self._schedule(k)
            for j in range(k + 1, k + 1 + self._depth):
                self._schedule(j)
            fut = self._pending.pop(k)
            for j in [j for j in self._pending if j < k]:       # skipped items: drop, do not hold their memory
                self._pending.pop(j).cancel()
        return fut.result()                                     # re-raises the reader's exception here

    def close(self):
        with self._lock:
            for f in self._pending.values():
                f.cancel()
            self._pending.clear()
        self._pool.shutdown(wait=True)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


class OrderedSink:
    """Runs ``submit(fn, *args)`` jobs one after another on a worker thread, in submission order.  ``submit`` blocks
    while ``depth`` jobs are outstanding (bounds the pinned buffers a caller has to keep alive) and re-raises the
    first failure of an earlier job; ``close()`` drains the queue and does the same."""

    def __init__(self, depth=4):
        self._pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="f2f-write")
        self._inflight = deque()
        self._depth = max(1, int(depth))

    def _reap(self, block_to):
        while self._inflight and (len(self._inflight) > block_to or self._inflight[0].done()):
            self._inflight.popleft().result()

    def submit(self, fn, *args, **kw):
        self._reap(self._depth - 1)
        self._inflight.append(self._pool.submit(fn, *args, **kw))

    def close(self):
        try:
            self._reap(0)
        finally:
            self._pool.shutdown(wait=True)

    def __enter__(self):
        return self

    def __exit__(self, et, ev, tb):
        if et is None:
            self.close()
        else:                                                   # already failing: do not mask the first error
            self._pool.shutdown(wait=True, cancel_futures=True)
        return False
