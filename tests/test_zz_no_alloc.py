"""gfe_filter_alignments_static: filter mode for hosts that must keep this library away from the CUDA
allocator.

Two kinds of caller are meant: a process whose HBM is already spoken for (a framework's caching
allocator or a memory manager took it at start-up, so a cudaMalloc from inside the library would fail
or upset their bookkeeping), and a service with other streams running on the GPU, which a cudaMalloc /
cudaFree would stall.  They hold the alignment table in device memory of their own; the entry point
takes device pointers only and keeps its two bitmaps in a fixed array built into the library.

Checked here, through the C ABI, against gfe_filter_alignments (the engine entry point, which does
allocate) on the same table — byte for byte, as every filter test in this suite is:
  * two calls in a row with DIFFERENT node sets, the second a subset of the first: the array still
    holds the first call's node set, so a missing clear would show as rows kept that should be dropped;
  * a smaller table after a larger one (stale per-step words beyond the new range must not matter);
  * the free device memory reported by the driver does not change across the later calls;
  * requests over the built-in limits are refused with GFE_ERR_LIMIT and a message, nothing enqueued.

The file sorts last on purpose: with `pytest -x` a failure here cannot hide another test.  It builds
its own engine, so it gives the same verdict alone in a fresh process and at the end of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GFE_ERR_LIMIT = -3
MAX_STEPS = 1 << 26     # GFE_STATIC_FILTER_MAX_STEPS
MAX_NODES = 1 << 20     # GFE_STATIC_FILTER_MAX_NODES


def _table(rng, A, n_nodes, lens):
    ln = rng.choice(lens, A).astype(np.uint32)
    off = np.zeros(A + 1, dtype=np.uint32)
    off[1:] = np.cumsum(ln)
    # most steps from the low half of the node space, so that half-sized node sets keep some rows
    node = np.where(rng.random(int(off[-1])) < 0.9, rng.integers(0, max(n_nodes // 2, 1), int(off[-1])),
                    rng.integers(0, n_nodes, int(off[-1])))
    steps = (node * 2 + rng.integers(0, 2, int(off[-1]))).astype(np.uint32)
    return off, steps


def _static_filter(lib, torch, d_off, d_steps, A, S, n_nodes, nodes, min_nodes):
    """One call of the no-allocation entry point; every array the call sees is a device array."""
    dev = d_off.device
    d_nodes = torch.from_numpy(np.ascontiguousarray(nodes, dtype=np.uint32).view(np.int32)).to(dev)
    d_keep = torch.full((max(A, 1),), 7, dtype=torch.uint8, device=dev)      # 7: "never written"
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    rc = lib.gfe_filter_alignments_static(d_off.data_ptr(), d_steps.data_ptr() if S else None, A, S, n_nodes,
                                          d_nodes.data_ptr() if d_nodes.numel() else None, d_nodes.numel(),
                                          min_nodes, d_keep.data_ptr(),
                                          torch.cuda.current_stream(dev).cuda_stream, err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(dev)[0]
    return d_keep[:A].cpu().numpy().astype(bool), free_before - free_after


def test_static_filter_equals_engine_filter_on_consecutive_calls():
    import torch
    import gfalign_b200
    lib = gfalign_b200.load_library()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)                 # the entry point works on the current device
    rng = np.random.default_rng(2026)
    n_nodes = 3000
    big = _table(rng, 60000, n_nodes, [0, 1, 2, 3, 4, 6, 9, 17, 40, 300])
    small = _table(rng, 777, 65, [0, 1, 2, 5, 33])
    wide = np.arange(0, n_nodes // 2, dtype=np.uint32)                      # first call: a large set
    narrow = wide[rng.random(wide.size) < 0.97]                             # second call: a subset of it
    leaked = []
    with gfalign_b200.Engine(0) as eng:
        for (off, steps), nn, requests in (
                (big, n_nodes, [(wide, 0), (narrow, 0), (narrow, 3), (np.zeros(0, np.uint32), 0), (wide, -1)]),
                (small, 65, [(np.arange(0, 40, dtype=np.uint32), 1), (np.arange(0, 20, dtype=np.uint32), 1)])):
            A, S = off.size - 1, int(off[-1])
            eng.set_alignments(off, steps, nn)
            d_off = torch.from_numpy(off.view(np.int32)).to(dev)
            d_steps = torch.from_numpy(steps.view(np.int32)).to(dev)
            kept_any = dropped_any = False
            for nodes, mn in requests:
                want = eng.filter_alignments(nodes, mn)                      # existing entry point
                got, grew = _static_filter(lib, torch, d_off, d_steps, A, S, nn, nodes, mn)
                assert np.array_equal(got, want), (A, nodes.size, mn, int(got.sum()), int(want.sum()))
                kept_any |= bool(want.any())
                dropped_any |= not bool(want.all())
                leaked.append(grew)
            assert kept_any and dropped_any                                  # the comparison saw both verdicts
    # the very first call may load the library's kernels onto the GPU; no later call takes device memory
    assert all(g <= 0 for g in leaked[1:]), leaked


def test_static_filter_rejects_requests_beyond_the_built_in_scratch():
    import torch
    import gfalign_b200
    lib = gfalign_b200.load_library()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    d = torch.zeros(16, dtype=torch.int32, device=dev)
    keep = torch.full((16,), 7, dtype=torch.uint8, device=dev)
    err = C.create_string_buffer(256)
    for S, nn, word in ((MAX_STEPS + 1, 10, b"steps"), (8, MAX_NODES + 1, b"nodes")):
        rc = lib.gfe_filter_alignments_static(d.data_ptr(), d.data_ptr(), 4, S, nn, d.data_ptr(), 2, 0,
                                              keep.data_ptr(), None, err, len(err))
        assert rc == GFE_ERR_LIMIT
        assert word in err.value and b"gfe_filter_alignments" in err.value
    torch.cuda.synchronize()
    assert bool((keep == 7).all())             # a refused request enqueues nothing
    # an empty table is not an error and touches nothing
    assert lib.gfe_filter_alignments_static(None, None, 0, 0, 5, None, 0, 0, None, None, err, len(err)) == 0
