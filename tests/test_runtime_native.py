"""Native host runtime (libpdb200_runtime.so): TCP store, watchdog, bucket assignment, flag mirror."""
import ctypes
import threading
import time

import pytest

from paddle_b200 import lib


def test_runtime_library_builds_and_loads():
    assert lib.runtime.available(), lib.runtime._err


def test_tcp_store_set_get_add_wait():
    from paddle_b200.distributed.store import TCPStore
    master = TCPStore("127.0.0.1", 0, is_master=True, world_size=2)
    client = TCPStore("127.0.0.1", master.port, is_master=False, world_size=2)
    master.set("k", b"hello")
    assert client.get("k") == b"hello" and client.check("k") and not client.check("nope")
    assert client.add("cnt", 2) == 2 and master.add("cnt", 3) == 5
    got = {}

    def waiter():
        got["v"] = client.get("late", timeout=10)
    t = threading.Thread(target=waiter)
    t.start()
    time.sleep(0.2)
    master.set("late", b"x" * 100000)                # larger than the first receive buffer
    t.join(10)
    assert got["v"] == b"x" * 100000
    with pytest.raises(TimeoutError):
        client.wait("never", timeout=0.2)
    assert master.delete_key("k") and not master.check("k")
    client.close()
    master.close()


def test_watchdog_flags_timeouts():
    rt = lib.runtime.require()
    assert rt.pdb200_wd_start(0, 100, 50) == 0
    rt.pdb200_wd_task_begin(7, 1, b"all_reduce", 1024, 1)
    rt.pdb200_wd_task_begin(7, 2, b"broadcast", 8, 0)
    rt.pdb200_wd_task_end(7, 1)
    assert rt.pdb200_wd_inflight() == 1
    time.sleep(0.4)
    assert rt.pdb200_wd_timeouts() >= 1
    buf = ctypes.create_string_buffer(4096)
    n = rt.pdb200_wd_dump(buf, 4096)
    text = buf.raw[:n].decode()
    assert "op=broadcast" in text and "[last-done]" in text and "op=all_reduce" in text
    rt.pdb200_wd_task_end(7, 2)
    assert rt.pdb200_wd_stop() == 0


def test_flag_mirror():
    rt = lib.runtime.require()
    rt.pdb200_flag_set(b"use_fused_comm", b"1")
    buf = ctypes.create_string_buffer(16)
    assert rt.pdb200_flag_get(b"use_fused_comm", buf, 16) == 1 and buf.value == b"1"
    assert rt.pdb200_flag_get(b"missing", buf, 16) == -1


@pytest.mark.gpu
def test_watchdog_reads_task_state_from_cuda_events():
    """Device truth: a task stays in flight until the GPU passes its end event (nobody calls wait()), a task queued
    behind a long kernel reads "not started", and the trace says which."""
    import ctypes
    import time
    import torch
    rt = lib.runtime.require()
    st = torch.cuda.current_stream().cuda_stream
    assert rt.pdb200_wd_start(0, 60000, 50) == 0
    try:
        torch.cuda.synchronize()
        torch.cuda._sleep(int(2e9))                                   # ~1 s of GPU time ahead of both tasks
        assert rt.pdb200_wd_task_begin_ev(3, 1, b"all_reduce", 4096, 0, st) == 1
        torch.cuda._sleep(int(2e8))
        assert rt.pdb200_wd_task_end_ev(3, 1, st) == 1
        assert rt.pdb200_wd_inflight() == 1                            # enqueued, not finished on the device
        buf = ctypes.create_string_buffer(4096)
        rt.pdb200_wd_dump(buf, 4096)
        assert b"state=not started" in buf.value, buf.value
        torch.cuda.synchronize()
        rt.pdb200_wd_sweep()
        assert rt.pdb200_wd_inflight() == 0                            # retired by the end event, without any wait()
        rt.pdb200_wd_dump(buf, 4096)
        assert b"[last-done] rank=0 group=3 seq=1" in buf.value
        p = rt.pdb200_wd_progress_ptr()
        assert p != 0
    finally:
        rt.pdb200_wd_stop()


def test_arena_heap_best_fit_coalesce_and_determinism():
    """csrc/runtime/heap.cc: the symmetric arena's offset heap.  Offsets depend on the call sequence only."""
    rt = lib.runtime.require()

    def run():
        h = rt.pdb200_heap_create(1 << 20, 4096)
        offs = [rt.pdb200_heap_alloc(h, n, 256) for n in (1000, 70000, 300, 200000, 5000)]
        assert all(o >= 4096 and o % 256 == 0 for o in offs)
        spans = sorted(zip(offs, (1000, 70000, 300, 200000, 5000)))
        assert all(a + n <= b for (a, n), (b, _) in zip(spans, spans[1:])), "allocations overlap"
        # free two neighbours -> one coalesced hole that takes a request neither piece could hold alone
        assert rt.pdb200_heap_free(h, offs[1]) == 0 and rt.pdb200_heap_free(h, offs[2]) == 0
        assert rt.pdb200_heap_free(h, offs[2]) == -1, "double free must be refused"
        big = rt.pdb200_heap_alloc(h, 70200, 256)
        assert big == offs[1], "the coalesced hole is the best fit"
        # best fit: a small request goes to the smallest hole that holds it, not to the big tail
        rt.pdb200_heap_free(h, offs[4])
        small = rt.pdb200_heap_alloc(h, 4000, 256)
        assert small == offs[4]
        assert rt.pdb200_heap_alloc(h, 1 << 21, 256) == -1
        seq = offs + [big, small]
        stats = [rt.pdb200_heap_stat(h, i) for i in range(7)]
        assert stats[6] == 1 << 20 and stats[3] <= stats[4] and stats[0] + stats[3] <= (1 << 20) - 4096
        rt.pdb200_heap_destroy(h)
        return seq, stats

    assert run() == run()


def test_reducer_core_releases_buckets_in_order_and_reports_unused():
    """csrc/runtime/reducer.cc: gradients arrive in any order, buckets are released strictly in bucket order."""
    import ctypes
    rt = lib.runtime.require()
    # 5 parameters: bucket 0 = {0, 1}, bucket 1 = {2} (sparse), bucket 2 = {3, 4}
    bucket_of = (ctypes.c_int32 * 5)(0, 0, 1, 2, 2)
    sparse = (ctypes.c_uint8 * 3)(0, 1, 0)
    h = rt.pdb200_reducer_create(ctypes.cast(bucket_of, ctypes.c_void_p), 5, ctypes.cast(sparse, ctypes.c_void_p), 3)
    assert h
    out = (ctypes.c_int32 * 3)()
    miss = (ctypes.c_int32 * 5)()
    nl = ctypes.c_int32(0)

    def mark(p):
        n = rt.pdb200_reducer_mark_ready(h, p, ctypes.cast(out, ctypes.c_void_p), 3)
        return n if n < 0 else list(out)[:n]

    assert mark(0) == -1                                  # not armed yet
    rt.pdb200_reducer_prepare(h)
    assert mark(4) == [] and mark(3) == []                # bucket 2 is full but buckets 0 and 1 come first
    assert mark(2) == []                                  # the sparse bucket is ready at once, still behind bucket 0
    assert mark(0) == [] and mark(0) == -1                # second mark of the same parameter is refused
    assert mark(1) == [0, 1, 2]                           # everything is released, in order
    assert rt.pdb200_reducer_finalize(h, 0, ctypes.cast(miss, ctypes.c_void_p), 5, ctypes.cast(out, ctypes.c_void_p), 3,
                                      ctypes.byref(nl)) == 0 and nl.value == 0
    # an iteration where parameter 1 stays unused
    rt.pdb200_reducer_prepare(h)
    assert mark(0) == [] and mark(3) == [] and mark(4) == [] and mark(2) == []
    n = rt.pdb200_reducer_finalize(h, 0, ctypes.cast(miss, ctypes.c_void_p), 5, ctypes.cast(out, ctypes.c_void_p), 3, ctypes.byref(nl))
    assert n == 1 and miss[0] == 1 and nl.value == 0      # not tolerated: reported, nothing released
    rt.pdb200_reducer_prepare(h)
    for p in (0, 3, 4, 2):
        mark(p)
    n = rt.pdb200_reducer_finalize(h, 1, ctypes.cast(miss, ctypes.c_void_p), 5, ctypes.cast(out, ctypes.c_void_p), 3, ctypes.byref(nl))
    assert n == 1 and miss[0] == 1 and list(out)[:nl.value] == [0, 1, 2]
    used = (ctypes.c_int32 * 5)()
    rt.pdb200_reducer_used_bitmap(h, ctypes.cast(used, ctypes.c_void_p))
    assert list(used) == [1, 0, 1, 1, 1]
    rt.pdb200_reducer_destroy(h)
