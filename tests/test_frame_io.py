"""Host-thread frame prefetcher / ordered writer (frame2frame_b200/frame_io.py): same results and
order as the reference's synchronous reads and writes (blind_denoising.py:170-203,233-253)."""
import threading
import time

import pytest

from frame2frame_b200.frame_io import OrderedSink, Prefetcher


def test_prefetcher_returns_every_item_and_reads_ahead():
    started, lock = [], threading.Lock()

    def read(k):
        with lock:
            started.append(k)
        time.sleep(0.01)
        return k * k

    with Prefetcher(read, 20, depth=3, workers=2) as seq:
        assert len(seq) == 20
        assert seq[0] == 0
        time.sleep(0.08)
        with lock:
            assert set(started) == {0, 1, 2, 3}          # depth items beyond the one asked for, no more
        assert [seq[k] for k in range(1, 20)] == [k * k for k in range(1, 20)]
        assert sorted(started) == list(range(20))        # nothing read twice
        assert seq[-1] == 361                            # random access still works (a fresh read)
        with pytest.raises(IndexError):
            seq[20]


def test_prefetcher_window_access_like_run_sequence():
    """run_sequence takes frames[0], then windows of `lookahead` consecutive frames."""
    reads = []
    with Prefetcher(lambda k: reads.append(k) or k, 11, depth=6, workers=3) as seq:
        got = [seq[0]]
        i = 1
        while i < 11:
            L = min(4, 11 - i)
            got += [seq[k] for k in range(i, i + L)]
            i += L
    assert got == list(range(11)) and sorted(reads) == list(range(11))


def test_prefetcher_raises_the_readers_error_at_the_item():
    def read(k):
        if k == 2:
            raise FileNotFoundError("frame %03d.png" % k)
        return k

    with Prefetcher(read, 5, depth=4) as seq:
        assert seq[0] == 0 and seq[1] == 1
        with pytest.raises(FileNotFoundError, match="002"):
            seq[2]
        assert seq[3] == 3


def test_ordered_sink_keeps_submission_order_and_bounds_the_queue():
    done, running = [], []

    def job(i):
        running.append(i)
        time.sleep(0.02 if i % 2 == 0 else 0.0)
        done.append(i)

    with OrderedSink(depth=3) as sink:
        for i in range(10):
            sink.submit(job, i)
            assert len(running) >= i - 2                 # at most `depth` jobs not yet started or finished
    assert done == list(range(10))


def test_ordered_sink_surfaces_a_failed_job():
    def job(i):
        if i == 1:
            raise OSError("disk full")

    sink = OrderedSink(depth=2)
    sink.submit(job, 0)
    sink.submit(job, 1)
    with pytest.raises(OSError, match="disk full"):
        for i in range(2, 6):
            sink.submit(job, i)
            time.sleep(0.01)
        sink.close()
    sink._pool.shutdown(wait=True)


def test_ordered_sink_does_not_mask_the_callers_exception():
    with pytest.raises(KeyError):
        with OrderedSink() as sink:
            sink.submit(time.sleep, 0.01)
            raise KeyError("loop failed")
