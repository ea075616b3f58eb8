"""kb200_blob_to_commitment_device_noalloc (include/kateth_b200.h): the allocation-free commit entry point, against the existing
kb200_blob_to_commitment_device on the same device buffers -- bit-exact (the bar of every test here), on two calls in a row that share the
static scratch array.  No counterpart in kateth; the file is named so that pytest collects it last."""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
BLOB = 131072


def _blobs(n, seed):
    """n canonical blobs: random bytes with the top byte of every element below 0x73 (r = 0x73ed...)."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(n, 4096, 32), dtype=np.uint8)
    a[:, :, 0] = rng.integers(0, 0x73, size=(n, 4096), dtype=np.uint8)
    return a.reshape(n, BLOB)


def _both(s, host_blobs):
    """-> (existing entry point's bytes and statuses, the no-alloc one's, device memory taken by the no-alloc call)."""
    import torch
    n = host_blobs.shape[0]
    d_b = torch.from_numpy(host_blobs).cuda()
    ref = torch.full((n, 48), 0xEE, dtype=torch.uint8, device="cuda")
    ref_st = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    out = torch.full((n, 48), 0xDD, dtype=torch.uint8, device="cuda")
    st = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc_ref = s.blob_to_commitment_device(d_b.data_ptr(), n, ref.data_ptr(), ref_st.data_ptr())
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = s.blob_to_commitment_device_noalloc(d_b.data_ptr(), n, out.data_ptr(), st.data_ptr())
    assert rc == 0, rc                                   # enqueue-only: errors of single blobs are in the status array
    torch.cuda.synchronize()
    taken = free0 - torch.cuda.mem_get_info()[0]
    return rc_ref, ref.cpu().numpy(), ref_st.cpu().numpy(), out.cpu().numpy(), st.cpu().numpy(), taken


def test_no_alloc_commit_equals_commit_on_two_calls_in_a_row(any_engine, oracle_setup):
    import kateth_b200 as K
    s = any_engine
    assert K.setup.NOALLOC_MAX_BLOBS == 64
    # call 1: the maximum batch (16 slices per blob); call 2: three blobs over 32 slices each, one of them invalid -- whatever call 1
    # left in the shared array must not show
    b1 = _blobs(64, 11)
    rc_ref, ref, ref_st, out, st, taken = _both(s, b1)
    assert rc_ref == 0 and not ref_st.any()
    assert not st.any()
    assert np.array_equal(out, ref)
    assert taken <= 0, "the no-alloc entry point took %d bytes of device memory" % taken
    assert bytes(out[0]) == oracle_setup.blob_to_commitment(bytes(b1[0]))

    b2 = _blobs(3, 12)
    b2[1, 32 * 7:32 * 8] = 0xFF                           # element 7 of blob 1 is not canonical
    rc_ref, ref, ref_st, out, st, taken = _both(s, b2)
    assert rc_ref == 2 and list(ref_st) == [0, 2, 0]
    assert list(st) == [0, 2, 0]
    assert np.array_equal(out, ref) and not out[1].any() and out[0].any() and out[2].any()
    assert taken <= 0

    # the same three blobs once more: same bytes (the array keeps call 2's contents; nothing of it may leak into call 3)
    _, _, _, out3, st3, _ = _both(s, b2)
    assert np.array_equal(out3, out) and list(st3) == [0, 2, 0]


def test_no_alloc_commit_rejects_batches_beyond_the_static_scratch(any_engine):
    import torch
    import kateth_b200 as K
    s = any_engine
    d = torch.zeros(16, dtype=torch.uint8, device="cuda")
    rc = s.blob_to_commitment_device_noalloc(d.data_ptr(), 65, d.data_ptr(), d.data_ptr())
    assert rc == -20                                      # KB200_E_ARG
    assert b"KB200_NOALLOC_MAX_BLOBS" in K.lib().kb200_last_error()
    assert s.blob_to_commitment_device_noalloc(0, 0, 0, 0) == 0
