"""The allocation-free entry point ``gpsat_local_select_count_noalloc`` (no counterpart in the reference)
against the existing ``gpsat_local_select``, through the C ABI.

Collected last on purpose (file name).  Passes alone in a fresh process and
after the other GPU tests (the library's fixed scratch keeps whatever an earlier call left in it).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _problem(seed, N, E):
    rng = np.random.default_rng(seed)
    # point table columns: x, y, t, obs; expert locations: x, y, t
    table = np.stack([rng.uniform(-500e3, 500e3, N), rng.uniform(-500e3, 500e3, N),
                      rng.integers(0, 30, N).astype(float), rng.standard_normal(N)], axis=1)
    locs = np.stack([rng.uniform(-400e3, 400e3, E), rng.uniform(-400e3, 400e3, E),
                     rng.integers(5, 25, E).astype(float)], axis=1)
    # the example config's local_select: t within +-4 days, (x, y) within 300 km (configs/example_local_expert_oi.json)
    conds = [(0, "<=", [2], [2], 4.0), (0, ">=", [2], [2], -4.0), (1, "<", [0, 1], [0, 1], 300e3)]
    return table, locs, conds


def test_noalloc_count_equals_allocating_entry_point_two_calls_in_a_row(engine):
    import torch
    from gpsat_b200 import local_experts as le

    dev = f"cuda:{engine.device}"
    # call 1 fills a large part of the scratch; call 2 is smaller, so stale counters of call 1 sit right behind the
    # ones it uses, and call 3 repeats call 1 after them
    problems = [_problem(1, 70_000, 300), _problem(2, 9_000, 40), _problem(1, 70_000, 300)]
    expected = [le.select_all(engine, t, l, c)[0] for t, l, c in problems]
    assert all(e.sum() > 0 for e in expected) and not np.array_equal(expected[0][:40], expected[1])
    inputs = [(torch.as_tensor(t, device=dev).contiguous(), torch.as_tensor(l, device=dev).contiguous(),
               torch.full((len(l),), -7, dtype=torch.int64, device=dev)) for t, l, _ in problems]
    torch.cuda.synchronize()
    for (tt, lt, ct), (_, _, conds), want in zip(inputs, problems, expected):
        before = torch.cuda.memory_allocated(dev)
        le.select_count_noalloc(engine, tt, lt, conds, ct)
        assert torch.cuda.memory_allocated(dev) == before
        np.testing.assert_array_equal(ct.cpu().numpy(), want)          # integer work: bit-exact


def test_noalloc_rejects_what_it_cannot_do(engine):
    import ctypes as C
    import torch
    from gpsat_b200 import _lib, local_experts as le

    dev = f"cuda:{engine.device}"
    table, locs, conds = _problem(3, 2_000, 8)
    tt, lt = torch.as_tensor(table, device=dev), torch.as_tensor(locs, device=dev)
    ct = torch.zeros(len(locs), dtype=torch.int64, device=dev)
    # beyond the fixed scratch: E * ceil(N / 1024) > 2^20 (sizes only; the buffers are never touched)
    arr = (_lib.SelectCond * 1)()
    rc = engine._lib.gpsat_local_select_count_noalloc(engine._h, 1024 * 4096, 4, C.c_void_p(tt.data_ptr()), 300, 3,
                                                      C.c_void_p(lt.data_ptr()), 0, arr, C.c_void_p(ct.data_ptr()))
    assert rc == 2 and "GPSAT_NOALLOC_MAX_CHUNK_COUNTS" in _lib.last_error()
    # host pointers are refused instead of being staged through an allocation
    rc = engine._lib.gpsat_local_select_count_noalloc(engine._h, len(table), 4, C.c_void_p(table.ctypes.data), len(locs),
                                                      3, C.c_void_p(lt.data_ptr()), 0, arr, C.c_void_p(ct.data_ptr()))
    assert rc == 2 and "device pointer" in _lib.last_error()
    # and the entry point still works afterwards
    le.select_count_noalloc(engine, tt.contiguous(), lt.contiguous(), conds, ct)
    np.testing.assert_array_equal(ct.cpu().numpy(), le.select_all(engine, table, locs, conds)[0])
