"""vo_match_static: the matcher for callers that may not allocate GPU memory.

Some users of libvofront.so cannot let a library call cudaMalloc: their device memory comes out of one arena with a
fixed budget per module, or they sit in a loop with a latency bound where cudaMalloc / cudaFree (which synchronise the
device) are banned, or they are a plug-in that only receives device pointers and a stream.  vo_create() is out for them
(it allocates every pool of the context), so include/vofront.h offers one context-free entry point whose scratch is an
array compiled into the library.  These tests call it through the C ABI with device buffers that the TEST allocates
(torch stands in for the caller's arena) and compare its match list with the existing entry point's, vo_match_host on a
context - bit for bit, the tolerance every matcher test in this suite uses.

The file is named so that pytest collects it last: with `pytest -x` a failure here cannot hide another test.  It uses
no state of the other tests, so it passes alone in a fresh process and after the rest of the suite alike.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATIC_MAX = 8192  # VO_STATIC_MATCH_MAX of include/vofront.h


def _descriptor_sets(rng, n1, n2, planted):
    """Packed descriptors (n x 16 uint32) with `planted` near-duplicates, so that the ratio test accepts some pairs."""
    d1 = rng.integers(0, 2**32, (n1, 16), dtype=np.uint64).astype(np.uint32)
    d2 = rng.integers(0, 2**32, (n2, 16), dtype=np.uint64).astype(np.uint32)
    k = min(planted, n1, n2)
    if k:
        src = rng.permutation(n1)[:k]
        dst = rng.permutation(n2)[:k]
        d2[dst] = d1[src]
        d2[dst, rng.integers(0, 16, k)] ^= (np.uint32(1) << rng.integers(0, 32, k).astype(np.uint32))  # one flipped test each
    return d1, d2


class _Caller:
    """A caller that owns its device memory: buffers from torch, handed to the C ABI as plain pointers."""

    def __init__(self):
        import torch
        from acs_visual_odometry_b200 import load_library
        self.torch, self.lib = torch, load_library()
        self.dev = torch.device("cuda", 0)
        torch.cuda.set_device(0)
        # allocated once, at "start-up", for the maximum the entry point takes
        self.q = torch.empty((STATIC_MAX, 16), dtype=torch.int32, device=self.dev)
        self.t = torch.empty((STATIC_MAX, 16), dtype=torch.int32, device=self.dev)
        self.pairs = torch.empty((STATIC_MAX, 2), dtype=torch.int32, device=self.dev)
        self.count = torch.empty((1,), dtype=torch.int32, device=self.dev)

    def match(self, d1, d2, ratio, bits):
        torch = self.torch
        n1, n2 = len(d1), len(d2)
        if n1:
            self.q[:n1].copy_(torch.from_numpy(d1.view(np.int32)))
        if n2:
            self.t[:n2].copy_(torch.from_numpy(d2.view(np.int32)))
        self.pairs.fill_(-7)
        self.count.fill_(-7)
        rc = self.lib.vo_match_static(C.c_void_p(self.q.data_ptr()), n1, C.c_void_p(self.t.data_ptr()), n2, ratio, bits,
                                      C.c_void_p(self.pairs.data_ptr()), C.c_void_p(self.count.data_ptr()), None)
        assert rc == 0, self.lib.vo_last_error(None)
        torch.cuda.synchronize()
        n = int(self.count.cpu()[0])
        got = self.pairs.cpu().numpy()
        assert 0 <= n <= n1 and (got[n:] == -7).all(), "wrote past the reported count"
        return got[:n].copy()


@pytest.fixture(scope="module")
def caller():
    return _Caller()


@pytest.fixture(scope="module")
def vo_ctx():
    """The existing entry point lives on a context; small frames, the library's largest keypoint count."""
    from acs_visual_odometry_b200 import VoFront
    ctx = VoFront(max_rows=64, max_cols=64, max_keypoints=STATIC_MAX, n_slots=2)
    yield ctx
    ctx.close()


def test_static_matcher_equals_context_matcher_on_two_calls_in_a_row(caller, vo_ctx):
    """Two calls back to back on the shared scratch, the second one SMALLER than the first so that the array still
    holds the first call's keys beyond the second call's queries and inside its slices; then the first shape again.
    Each must equal vo_match_host exactly, in both distance modes."""
    from acs_visual_odometry_b200 import MATCH_FULL512, MATCH_REF32
    rng = np.random.default_rng(2026)
    shapes = [(2048, 1999, 600), (700, 301, 200), (2048, 1999, 600)]
    sets = [_descriptor_sets(rng, *s) for s in shapes]
    for bits in (MATCH_REF32, MATCH_FULL512):
        for (d1, d2), shape in zip(sets, shapes):
            got = caller.match(d1, d2, 0.75, bits)
            want = vo_ctx.match_host(d1, d2, 0.75, bits)
            assert got.shape == want.shape and np.array_equal(got, want), f"bits={bits} shape={shape}: {len(got)} vs {len(want)}"
            if bits == MATCH_FULL512:
                assert len(want) >= shape[2] // 2, "planted duplicates should be accepted"
    # alternating modes on the same array, different ratios
    for ratio in (0.5, 1.0):
        for bits in (MATCH_FULL512, MATCH_REF32):
            d1, d2 = sets[1]
            assert np.array_equal(caller.match(d1, d2, ratio, bits), vo_ctx.match_host(d1, d2, ratio, bits)), (ratio, bits)


def test_static_matcher_at_its_maximum_and_on_ragged_and_empty_inputs(caller, vo_ctx):
    from acs_visual_odometry_b200 import MATCH_FULL512, MATCH_REF32
    rng = np.random.default_rng(77)
    d1, d2 = _descriptor_sets(rng, STATIC_MAX, STATIC_MAX, 3000)
    for bits in (MATCH_REF32, MATCH_FULL512):
        assert np.array_equal(caller.match(d1, d2, 0.75, bits), vo_ctx.match_host(d1, d2, 0.75, bits)), bits
    # ragged tails of the 128-wide tiles and of the 8 slices, fewer train rows than slices, one train row (no second best)
    for n1, n2 in ((1, 2), (129, 7), (5, 1), (127, 1025), (300, 8)):
        a, b = _descriptor_sets(rng, n1, n2, min(n1, n2))
        for bits in (MATCH_REF32, MATCH_FULL512):
            assert np.array_equal(caller.match(a, b, 0.75, bits), vo_ctx.match_host(a, b, 0.75, bits)), (n1, n2, bits)
    # empty sides give zero matches, like the reference (feature_matching_parallel.cpp:57)
    a, b = _descriptor_sets(rng, 10, 10, 5)
    assert len(caller.match(a[:0], b, 0.75, MATCH_REF32)) == 0
    assert len(caller.match(a, b[:0], 0.75, MATCH_FULL512)) == 0


def test_static_matcher_refuses_what_it_cannot_do(caller):
    """Sizes beyond the compiled-in scratch, host pointers and bad modes are errors with a message, never a launch."""
    lib = caller.lib
    q, t, p, c = (C.c_void_p(x.data_ptr()) for x in (caller.q, caller.t, caller.pairs, caller.count))
    assert lib.vo_match_static(q, STATIC_MAX + 1, t, 10, 0.75, 32, p, c, None) == -1
    msg = lib.vo_last_error(None).decode()
    assert "VO_STATIC_MATCH_MAX" in msg and str(STATIC_MAX) in msg and str(STATIC_MAX + 1) in msg, msg
    assert lib.vo_match_static(q, 10, t, STATIC_MAX + 1, 0.75, 32, p, c, None) == -1
    assert lib.vo_match_static(q, 10, t, 10, 0.75, 64, p, c, None) == -1
    assert lib.vo_match_static(q, -1, t, 10, 0.75, 32, p, c, None) == -1
    assert lib.vo_match_static(q, 10, t, 10, 0.75, 32, p, None, None) == -1
    host = np.zeros((10, 16), np.uint32)
    assert lib.vo_match_static(host.ctypes.data_as(C.c_void_p), 10, t, 10, 0.75, 32, p, c, None) == -1
    assert "device memory" in lib.vo_last_error(None).decode()
    # and the array is still usable afterwards
    rng = np.random.default_rng(5)
    a, b = _descriptor_sets(rng, 64, 64, 32)
    assert len(caller.match(a, b, 0.75, 512)) >= 16
