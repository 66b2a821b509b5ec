"""Host logic of the device-resident HARQ soft-buffer pool (srsran_project_23.5_b200/host/rx_softbuffer_pool_cuda.cpp)
against the reference's rx_softbuffer_pool_impl compiled from /root/reference, driven through the same random sequence of
reserve / unlock / release / run_slot operations (rx_softbuffer_pool_impl.cpp:27-76, rx_softbuffer_impl.h:34-262):
validity of every reservation, codeblock counts, persistence of CRC flags and of the stored message bytes (which follow the
first-fit codeblock entries of rx_softbuffer_codeblock_pool.h) must be identical. Runs on CPU: the product pool is created
without device memory. The GPU half of the pool (soft bits in the shared HBM arena) is covered by test_gpu_dropin.py."""
import ctypes
import os

import numpy as np
import pytest

import oracles as O

DROPIN = os.path.join(O.ORACLE_DIR, "_ref", "libdropin.so")


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(DROPIN):
        pytest.skip("oracle/_ref/libdropin.so was not built (no reference tree in the build container)")
    try:
        L = ctypes.CDLL(DROPIN)
    except OSError as e:
        pytest.skip("libdropin.so does not load here: %s" % e)
    L.ref_pool_create.restype = ctypes.c_void_p
    L.ref_pool_create.argtypes = [ctypes.c_char_p, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint]
    for name, args in (("ref_pool_destroy", [ctypes.c_void_p]),
                       ("ref_pool_reserve", [ctypes.c_void_p] + [ctypes.c_uint] * 4),
                       ("ref_pool_nof_codeblocks", [ctypes.c_void_p, ctypes.c_int]),
                       ("ref_pool_crc", [ctypes.c_void_p, ctypes.c_int, ctypes.c_uint, ctypes.c_int]),
                       ("ref_pool_data_byte", [ctypes.c_void_p, ctypes.c_int, ctypes.c_uint, ctypes.c_int]),
                       ("ref_pool_unlock", [ctypes.c_void_p, ctypes.c_int]),
                       ("ref_pool_release", [ctypes.c_void_p, ctypes.c_int]),
                       ("ref_pool_run_slot", [ctypes.c_void_p, ctypes.c_uint]),
                       ("ref_pool_stats", [ctypes.c_void_p, ctypes.POINTER(ctypes.c_ulonglong)]),
                       ("ref_pool_time_reserve", [ctypes.c_void_p] + [ctypes.c_uint] * 3)):
        getattr(L, name).argtypes = args
    L.ref_pool_time_reserve.restype = ctypes.c_double
    return L


def _stats(lib, pool):
    out = (ctypes.c_ulonglong * 6)()
    assert lib.ref_pool_stats(pool, out) == 0
    return dict(zip(("reservations", "refused_in_use", "refused_no_buffer", "refused_no_slots", "expired", "released"),
                    [int(v) for v in out]))


@pytest.mark.parametrize("seed,nbuf,ncb,expire", [(1, 4, 8, 5), (2, 2, 6, 3), (3, 6, 40, 10), (4, 3, 5, 1)])
def test_pool_state_machine_matches_reference(lib, seed, nbuf, ncb, expire):
    rng = np.random.default_rng(seed)
    pools = [lib.ref_pool_create(b"sw", nbuf, ncb, expire), lib.ref_pool_create(b"cuda-host", nbuf, ncb, expire)]
    assert all(pools)
    held = {}  # token -> nof codeblocks
    slot = 0
    n_valid = n_invalid = 0
    for step in range(3000):
        op = rng.integers(0, 10)
        if op < 4:
            rnti, harq, n = int(rng.integers(1, 6)), int(rng.integers(0, 2)), int(rng.choice([1, 2, 3, 5, ncb]))
            toks = [lib.ref_pool_reserve(p, slot, rnti, harq, n) for p in pools]
            assert toks[0] == toks[1], (step, toks)
            if toks[0] < 0:
                n_invalid += 1
                continue
            n_valid += 1
            t = toks[0]
            assert [lib.ref_pool_nof_codeblocks(p, t) for p in pools] == [n, n]
            for cb in range(n):
                # (CRC flags of a freshly sized buffer are uninitialised in the reference - static_vector::resize - and
                # zeroed by the decoder on new data: their persistence is checked in the directed test below)
                a, b = [lib.ref_pool_data_byte(p, t, cb, -1) for p in pools]
                assert a == b, (step, cb)
                if rng.integers(0, 2):
                    v = int(rng.integers(0, 2))
                    w = int(rng.integers(0, 256))
                    for p in pools:
                        lib.ref_pool_crc(p, t, cb, v)
                        lib.ref_pool_data_byte(p, t, cb, w)
            held[t] = n
        elif op < 7 and held:
            t = int(rng.choice(list(held)))
            for p in pools:
                lib.ref_pool_unlock(p, t)
            del held[t]
        elif op < 8 and held:
            t = int(rng.choice(list(held)))
            for p in pools:
                lib.ref_pool_release(p, t)
            del held[t]
        else:
            slot += int(rng.integers(1, 3))
            for p in pools:
                lib.ref_pool_run_slot(p, slot)
    assert n_valid > 100 and n_invalid > 10
    for p in pools:
        lib.ref_pool_destroy(p)


def test_pool_takes_more_than_52_codeblocks(lib):
    p = lib.ref_pool_create(b"cuda-host", 2, 200, 10)
    t = lib.ref_pool_reserve(p, 0, 7, 1, 152)
    assert t >= 0 and lib.ref_pool_nof_codeblocks(p, t) == 152
    assert lib.ref_pool_reserve(p, 0, 8, 1, 152) < 0  # only 48 entries left: reservation fails like the reference's
    lib.ref_pool_destroy(p)


def test_crc_flags_persist_across_reservations_of_the_same_process(lib):
    for kind in (b"sw", b"cuda-host"):
        p = lib.ref_pool_create(kind, 3, 12, 4)
        t = lib.ref_pool_reserve(p, 0, 17, 2, 5)
        assert t >= 0
        for cb in range(5):
            lib.ref_pool_crc(p, t, cb, cb % 2)
        lib.ref_pool_unlock(p, t)
        lib.ref_pool_run_slot(p, 1)
        t = lib.ref_pool_reserve(p, 2, 17, 2, 5)  # retransmission: same process, same size -> same entries and flags
        assert t >= 0 and [lib.ref_pool_crc(p, t, cb, -1) for cb in range(5)] == [0, 1, 0, 1, 0], kind
        assert lib.ref_pool_reserve(p, 2, 17, 2, 5) < 0, kind  # locked by the holder above
        lib.ref_pool_unlock(p, t)
        lib.ref_pool_run_slot(p, 2 + 4)  # expired: entries freed
        t2 = lib.ref_pool_reserve(p, 7, 18, 0, 12)  # all 12 entries available again
        assert t2 >= 0, kind
        lib.ref_pool_release(p, t2)
        lib.ref_pool_run_slot(p, 8)
        assert lib.ref_pool_reserve(p, 8, 19, 0, 12) >= 0, kind
        lib.ref_pool_destroy(p)


def test_pool_counts_what_happens_to_reservations(lib):
    """The product pool records expiries, releases and refused reservations (the reference loses HARQ state silently)."""
    p = lib.ref_pool_create(b"cuda-host", 2, 10, 3)
    assert lib.ref_pool_stats(lib.ref_pool_create(b"sw", 2, 10, 3), (ctypes.c_ulonglong * 6)()) == -1
    a = lib.ref_pool_reserve(p, 0, 1, 0, 6)
    assert a >= 0
    assert lib.ref_pool_reserve(p, 0, 1, 0, 6) < 0      # same identifier, still held by `a`
    assert lib.ref_pool_reserve(p, 0, 2, 0, 6) < 0      # 4 entries left
    b = lib.ref_pool_reserve(p, 0, 3, 0, 4)
    assert b >= 0
    assert lib.ref_pool_reserve(p, 0, 4, 0, 1) < 0      # both buffers taken
    lib.ref_pool_unlock(p, a)
    lib.ref_pool_release(p, b)
    lib.ref_pool_run_slot(p, 1)                         # b's entries return, a stays until its deadline
    assert _stats(lib, p) == dict(reservations=2, refused_in_use=1, refused_no_buffer=1, refused_no_slots=1, expired=0,
                                  released=1)
    lib.ref_pool_run_slot(p, 3)                         # deadline of a = slot 0 + 3
    assert _stats(lib, p)["expired"] == 1
    lib.ref_pool_destroy(p)


def test_reserving_a_full_slot_buffer_in_a_large_pool_is_fast(lib):
    """152 codeblock entries behind 30 000 occupied ones in a 38 912-entry pool (16 cells x 16 HARQ x 152): the lowest-free
    bitmap makes the reservation independent of how full the pool is; the reference's first-fit scan walks the occupied
    entries for every codeblock."""
    ours = lib.ref_pool_create(b"cuda-host", 4, 38912, 10)
    ns = lib.ref_pool_time_reserve(ours, 30000, 152, 200)
    lib.ref_pool_destroy(ours)
    assert 0 < ns < 10000, ns
