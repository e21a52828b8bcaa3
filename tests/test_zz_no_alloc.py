"""b2_gemv_noalloc -- the allocation-free twin of the b2_gemv hook, for callers that cannot call
cudaMalloc (a server whose KV cache already fills the GPU, code inside someone else's memory
pool).  It keeps its device staging in a static array inside the library, shared by all calls,
so the things worth checking are: same bits as b2_gemv, nothing stale leaking from one call into
the next (second call is smaller and of another type than the first), and a clean refusal
beyond the array's size.  This file sorts last on purpose: if it fails under `-x`, every other
test has already run."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(pkg, t, rows, k, seed, with_gain):
    from gguf import GGMLQuantizationType as T
    rng = np.random.default_rng(seed)
    W = pkg.synth.random_blocks(T(t), rows, k, k ** -0.5, rng)
    x = rng.standard_normal(k).astype(np.float32)
    g = (1 + 0.02 * rng.standard_normal(k)).astype(np.float32) if with_gain else None
    return W, x, g


def test_two_calls_in_a_row_match_the_allocating_entry_point(pkg):
    E = pkg.engine
    # first call: large Q4_0 problem with the fused RMSNorm prologue, fills most of the scratch
    W1, x1, g1 = _case(pkg, E.TYPE_Q4_0, 1536, 4096, 1, True)
    want1 = E.gemv(E.TYPE_Q4_0, W1, 1536, 4096, x1, gain=g1, eps=1e-5)
    got1 = E.gemv_noalloc(E.TYPE_Q4_0, W1, 1536, 4096, x1, gain=g1, eps=1e-5)
    assert np.array_equal(got1.view(np.uint32), want1.view(np.uint32))
    # second call, straight after: smaller, different type and row stride, no gain -- whatever the
    # first call left in the shared array must not show
    W2, x2, _ = _case(pkg, E.TYPE_Q6_K, 66, 1280, 2, False)
    want2 = E.gemv(E.TYPE_Q6_K, W2, 66, 1280, x2)
    got2 = E.gemv_noalloc(E.TYPE_Q6_K, W2, 66, 1280, x2)
    assert np.array_equal(got2.view(np.uint32), want2.view(np.uint32))
    # and the first problem again, to see the second left nothing behind either
    again = E.gemv_noalloc(E.TYPE_Q4_0, W1, 1536, 4096, x1, gain=g1, eps=1e-5)
    assert np.array_equal(again.view(np.uint32), want1.view(np.uint32))


def test_requests_beyond_the_static_scratch_are_refused(pkg):
    E = pkg.engine
    rows, k = 4096, 4096                      # 9.4 MB of Q4_0 blocks alone: more than the 8 MiB array
    W = np.zeros(rows * k // 32 * 18, np.uint8)
    with pytest.raises(E.B2Error) as e:
        E.gemv_noalloc(E.TYPE_Q4_0, W, rows, k, np.zeros(k, np.float32))
    assert e.value.code == -6 and "static scratch" in str(e.value)
    # the refusal did no partial work: the next valid call is still right
    W2, x2, _ = _case(pkg, E.TYPE_Q4_0, 64, 512, 3, False)
    assert np.array_equal(E.gemv_noalloc(E.TYPE_Q4_0, W2, 64, 512, x2), E.gemv(E.TYPE_Q4_0, W2, 64, 512, x2))


def test_calls_leave_free_device_memory_unchanged(pkg):
    """The point of the entry: no device allocation.  After one call that has done the once-per-device
    set-up, further calls must leave cudaMemGetInfo's free byte count exactly where it was."""
    import torch
    E = pkg.engine
    W, x, g = _case(pkg, E.TYPE_Q4_0, 512, 2048, 4, True)
    E.gemv_noalloc(E.TYPE_Q4_0, W, 512, 2048, x, gain=g, eps=1e-5)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        E.gemv_noalloc(E.TYPE_Q4_0, W, 512, 2048, x, gain=g, eps=1e-5)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)


def test_absurd_sizes_are_refused_before_any_arithmetic_can_wrap(pkg):
    E = pkg.engine
    lib = E.lib()
    y = np.zeros(4, np.float32)
    W = np.zeros(64, np.uint8)
    x = np.zeros(32, np.float32)
    # rows * row_bytes would wrap a 64-bit size to something small; the bound on rows alone must catch it
    rc = lib.b2_gemv_noalloc(E.TYPE_Q4_0, W.ctypes.data, 2 ** 60, 32, x.ctypes.data, None, 0.0, y.ctypes.data)
    assert rc == -6, rc
