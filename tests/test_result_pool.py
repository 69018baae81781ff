"""The facade recycles the buffers of large result arrays (fresh pages cost more than the conversion that fills them); a
buffer must never be handed out again while any array or view of an earlier result is alive.  No GPU needed."""
import gc

import numpy as np


def _addr(a):
    return a.__array_interface__["data"][0]


def test_result_buffers_are_recycled_only_when_nothing_refers_to_them(pkg):
    pool = pkg.facade._ResultPool()
    shape = (1_000_000, 2)
    a = pool.empty(shape, np.float64)
    assert a.shape == shape and a.dtype == np.float64 and a.flags.c_contiguous and a.flags.writeable
    a[:] = 1.0
    addr = _addr(a)
    view = a[:, 1]
    del a
    gc.collect()
    b = pool.empty(shape, np.float64)                       # a view of the first result is alive: a different buffer
    b[:] = 2.0
    assert _addr(b) != addr and float(view.sum()) == 1_000_000.0
    del view
    gc.collect()
    c = pool.empty(shape, np.float64)                       # now nothing refers to it: the same pages again
    assert _addr(c) == addr
    nested = c.reshape(-1)[5:][::2]                          # a view of a view keeps it out as well
    del c
    gc.collect()
    d = pool.empty(shape, np.float64)
    assert _addr(d) != addr and nested.base is not None
    e = pool.empty(shape, np.float64)
    assert len({_addr(b), _addr(d), _addr(e)}) == 3          # live results never share memory
    small = pool.empty((10, 2), np.float64)                  # small results come from numpy's allocator
    assert small.shape == (10, 2)
    other = pool.empty((1_000_000,), np.uint8)
    assert other.dtype == np.uint8 and other.shape == (1_000_000,)


def test_long_lived_results_do_not_clog_the_pool(pkg):
    pool = pkg.facade._ResultPool()
    shape = (600_000, 2)
    keep = [pool.empty(shape, np.float64) for _ in range(5)]            # the caller holds on to five results
    assert len({_addr(a) for a in keep}) == 5
    x = pool.empty(shape, np.float64)
    addr = _addr(x)
    del x
    gc.collect()
    y = pool.empty(shape, np.float64)                                   # the released one is recycled all the same
    assert _addr(y) == addr and len(pool._free) <= pool.MAX_HELD
