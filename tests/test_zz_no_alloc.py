"""b200_ctc_viterbi_static: the Viterbi alignment with its back-pointer scratch in a static device array (no caller
work buffer, no allocation inside the call).  Checked through the C ABI against b200_ctc_viterbi on the same inputs,
bit for bit (alignments are integer work), on two calls in a row that share the static array.

Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(T, V, L, seed):
    g = torch.Generator().manual_seed(seed)
    lpz = torch.log_softmax(torch.randn(T, V, generator=g), dim=1).cuda().contiguous()
    y = torch.randint(1, V, (L,), generator=g)
    if L >= 2:
        y[1] = y[0]
    return lpz, y.cuda()


def _existing(lib, lpz, y):
    T, V = lpz.shape
    L = y.numel()
    nbytes = lib.b200_ctc_viterbi_work_bytes(T, L)
    work = torch.empty(max(nbytes, 2), dtype=torch.uint8, device="cuda")
    align = torch.full((T,), -7, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.b200_ctc_viterbi(lpz.data_ptr(), T, V, y.data_ptr(), L, 0, align.data_ptr(), work.data_ptr(), nbytes, st)
    assert rc == 0, lib.b200_last_error()
    return align


def _static(lib, lpz, y):
    T, V = lpz.shape
    align = torch.full((T,), -7, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.b200_ctc_viterbi_static(lpz.data_ptr(), T, V, y.data_ptr(), y.numel(), 0, align.data_ptr(), st)
    assert rc == 0, lib.b200_last_error()
    return align


def test_static_scratch_viterbi_matches_the_existing_entry_point(lib):
    assert lib.b200_ctc_viterbi_static_max_bytes() == 4 << 20
    # two calls in a row with different shapes: the second one finds the first one's back-pointers in the array
    for T, V, L, seed in ((300, 50, 40, 1), (120, 20, 7, 2), (300, 50, 40, 3)):
        lpz, y = _inputs(T, V, L, seed)
        want = _existing(lib, lpz, y)
        got = _static(lib, lpz, y)
        torch.cuda.synchronize()
        assert torch.equal(got, want)
        assert int(got.min()) >= 0 and int(got.max()) < V
    # the same input twice in a row gives the same alignment
    lpz, y = _inputs(200, 30, 25, 4)
    a, b = _static(lib, lpz, y), _static(lib, lpz, y)
    torch.cuda.synchronize()
    assert torch.equal(a, b) and torch.equal(a, _existing(lib, lpz, y))


def test_static_scratch_viterbi_rejects_sizes_beyond_the_array(lib):
    lpz, y = _inputs(8, 6, 2, 5)
    align = torch.zeros(8, dtype=torch.int64, device="cuda")
    # T * (2L+1) * 2 bytes > 4 MiB: refused before anything is launched (the pointers are never dereferenced)
    rc = lib.b200_ctc_viterbi_static(lpz.data_ptr(), 4000, 6, y.data_ptr(), 1000, 0, align.data_ptr(), None)
    assert rc == 3 and b"static scratch" in lib.b200_last_error()
