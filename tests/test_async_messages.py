"""Async messages (reference message/async_queue.hpp, async_writer.hpp, async_reader.hpp; tests/message/async_queue.cpp):
device MPMC rings — host and kernel producers, host and kernel consumers — against the oracle's queue.

Order contract (same as the reference's moodycamel queue): FIFO per producer, unspecified between concurrent producers.
Host-produced streams are therefore compared byte for byte; messages pushed by the warps of a kernel are compared as a
multiset, and what a drain does to the world is compared bit-exactly (integer atomics make it order-free)."""
import struct

import numpy as np
import pytest

import scenarios
from heliosengine_b200 import capi, worlds

ALERT = "helios::bench::LowHealthAlert"
EVENT = "helios::bench::EventMsg"
ALERT_DTYPE = np.dtype([("who", "<u8"), ("hp", "<i4"), ("pad", "<u4")])


def queue_kats(backend):
    """tests/message/async_queue.cpp through the C ABI: Enqueue / EnqueueBulk / Dequeue / Into(max_count) / Clear / SizeApprox."""
    w = capi.World(backend)
    out = {}
    w.async_message_register(EVENT, 4, capacity=1024)
    w.async_message_register(EVENT, 4, capacity=1024)  # idempotent (:401)
    out["empty"] = w.async_message_size(EVENT)
    out["dequeue_empty"] = w.async_message_dequeue(EVENT).size
    w.async_message_write(EVENT, np.array([33], dtype=np.int32))
    out["one"] = w.async_message_size(EVENT)
    out["dequeued"] = w.async_message_dequeue(EVENT, 1).view(np.int32).tolist()
    w.async_message_write(EVENT, np.array([], dtype=np.int32))  # bulk enqueue of an empty range (:134)
    out["after_empty_bulk"] = w.async_message_size(EVENT)
    w.async_message_write(EVENT, np.array([1, 2, 3], dtype=np.int32))
    out["into_max2"] = w.async_message_dequeue(EVENT, 2).view(np.int32).tolist()  # "Respects max_count limit" (:233)
    out["left"] = w.async_message_size(EVENT)
    w.async_message_write(EVENT, np.array([4, 5], dtype=np.int32))
    out["fifo"] = w.async_message_dequeue(EVENT).view(np.int32).tolist()
    w.async_message_write(EVENT, np.arange(10, dtype=np.int32))
    w.async_message_clear(EVENT)
    out["after_clear"] = w.async_message_size(EVENT)
    # ring wrap-around: 1024 slots, 700 in / 700 out three times
    seen = []
    for k in range(3):
        w.async_message_write(EVENT, np.arange(700, dtype=np.int32) + 1000 * k)
        seen.append(w.async_message_dequeue(EVENT).view(np.int32).copy())
    out["wrap_ok"] = all(np.array_equal(seen[k], np.arange(700, dtype=np.int32) + 1000 * k) for k in range(3))
    out["unregistered_size"] = w.async_message_size("helios::bench::Nope")  # HasAsyncMessages of an unknown type: false
    w.async_message_clear("helios::bench::Nope")  # no-op (:452)
    w.close()
    return out


EXPECTED_KATS = {"empty": 0, "dequeue_empty": 0, "one": 1, "dequeued": [33], "after_empty_bulk": 0, "into_max2": [1, 2],
                 "left": 1, "fifo": [3, 4, 5], "after_clear": 0, "wrap_ok": True, "unregistered_size": 0}


def alert_world(backend, n, frames, mode, partial_host_drain=False):
    """S0 regen -> S1 alert_low_health (kernel producer) -> host peeks -> S2 heal_alerted (kernel consumer)."""
    w = capi.World(backend)
    worlds.register_all(w, ["Position", "Health"])
    H = worlds.cname("Health")
    w.spawn_batch([worlds.cname("Position"), H], n, [worlds.column("Position", n, 5100), worlds.column("Health", n, 5101)],
                  want_entities=False)
    w.spawn_batch([H], n // 3, [worlds.column("Health", n // 3, 5102)], want_entities=False)
    w.async_message_register(ALERT, 16, capacity=2 * n)
    params = struct.pack("<QiIQ", capi.fnv1a(ALERT), 40, w.slots[H], 0)
    emit = w.schedule("Emit")
    emit.add_system("S0_regen", "regen", w.query([(H, True)]))
    emit.add_system("S1_alert", "alert_low_health", w.query([(H, False)]), params=params)
    emit.build()
    heal = w.schedule("Heal")
    heal.add_system("S2_heal", "heal_alerted", None, params=params, extra_writes=[H])
    heal.build()
    out = {}
    for f in range(frames):
        emit.run_and_wait(mode)
        out[f"frame{f}_queued"] = w.async_message_size(ALERT)
        if partial_host_drain:  # a host consumer takes some, the kernel consumer the rest
            taken = w.async_message_dequeue(ALERT, 5).view(ALERT_DTYPE)
            out[f"frame{f}_host_took"] = int(taken.size)
        else:  # drain everything on the host to look at it, then put it back (single producer: FIFO) for the kernel
            alerts = w.async_message_dequeue(ALERT).view(ALERT_DTYPE)
            order = np.lexsort((alerts["hp"], alerts["who"]))
            out[f"frame{f}_alerts_sorted"] = scenarios._bytes_digest(alerts[order])
            w.async_message_write(ALERT, alerts)
        heal.run_and_wait(mode)
        out[f"frame{f}_left"] = w.async_message_size(ALERT)
        if not partial_host_drain:
            out[f"frame{f}_world"] = scenarios.snapshot_digest(w.snapshot())
    w.close()
    return out


def test_oracle_async_queue_kats(oracle):
    assert queue_kats(oracle) == EXPECTED_KATS


@pytest.mark.gpu
def test_gpu_async_queue_kats(gpu):
    assert queue_kats(gpu) == EXPECTED_KATS


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [capi.EXEC_SERIAL, capi.EXEC_STREAMS, capi.EXEC_GRAPH])
def test_gpu_kernel_producers_and_consumers_match_oracle(gpu, oracle, mode):
    want = alert_world(oracle, 20011, 3, capi.EXEC_SERIAL)
    assert want["frame0_queued"] > 1000
    assert alert_world(gpu, 20011, 3, mode) == want


@pytest.mark.gpu
def test_gpu_host_and_kernel_consumers_share_a_ring(gpu, oracle):
    got, want = (alert_world(b, 4099, 2, capi.EXEC_SERIAL, partial_host_drain=True) for b in (gpu, oracle))
    # WHICH five alerts the host consumer gets is unspecified between concurrent producers (as in the reference), so
    # frame 1 sees up to five different entities still below the threshold; everything else is fixed
    for key in ("frame0_queued", "frame0_host_took", "frame0_left", "frame1_host_took", "frame1_left"):
        assert got[key] == want[key], key
    assert got["frame0_host_took"] == 5 and got["frame0_left"] == 0
    assert abs(got["frame1_queued"] - want["frame1_queued"]) <= 5


@pytest.mark.gpu
def test_gpu_ring_overflow_is_reported(gpu):
    w = capi.World(gpu)
    w.async_message_register(EVENT, 4, capacity=1024)
    w.async_message_write(EVENT, np.zeros(1000, dtype=np.int32))
    with pytest.raises(capi.HeliosError) as e:
        w.async_message_write(EVENT, np.zeros(100, dtype=np.int32))
    assert e.value.status == capi.ERR_CAPACITY
    assert w.async_message_size(EVENT) == 1000
    w.close()


def test_oracle_alert_pipeline_is_consistent(oracle):
    r = alert_world(oracle, 4099, 2, capi.EXEC_SERIAL)
    assert r["frame0_queued"] > 0 and r["frame0_left"] == 0 and r["frame1_left"] == 0
    assert r["frame1_queued"] < r["frame0_queued"]  # healed entities stop alerting
