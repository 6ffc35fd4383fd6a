"""cfd_upload_noalloc: the upload path for hosts that cannot allocate device memory after start-up (a driver
whose runtime owns all free HBM, or a loop that re-uploads boundary data every iteration and must not meet
cudaMalloc's device-wide synchronisation).  It stages through a fixed 32 MiB buffer inside the library image
instead of a cudaMalloc'ed one, so it must (a) land exactly the same device contents as cfd_upload, also on a
second call that reuses the shared buffer with a different, smaller grid, and (b) refuse grids beyond its limit
with a clear message.  Collected last on purpose (file name), so a failure here cannot hide other tests."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_upload_matches_upload_on_two_calls_in_a_row(cfd):
    rng = np.random.default_rng(42)
    # call 1: a grid that nearly fills the staging buffer's row pattern; call 2: a smaller, differently shaped one
    for (nicv, njcv) in ((700, 900), (33, 130)):
        ref = cfd.CavityFVM(nicv, njcv)
        new = cfd.CavityFVM(nicv, njcv)
        a = np.asfortranarray(rng.standard_normal((ref.ni, ref.nj)))
        for name in ("u", "ae"):              # a plain field and one of the interleaved coefficient arrays
            ref.upload(name, a)
            new.upload_noalloc(name, a)
            got, want = new.download(name), ref.download(name)
            assert np.array_equal(got, want), name      # integer-exact data movement: same bits
            assert np.array_equal(got, a), name
        ref.close()
        new.close()


def test_noalloc_upload_honours_a_padded_leading_dimension(cfd):
    """A host array whose leading dimension is larger than ni (the reference's (it, jt) storage, control.f90:22-25):
    the pitch argument of the staged copy must skip the padding rows exactly as cfd_upload does."""
    rng = np.random.default_rng(7)
    ref, new = cfd.CavityFVM(61, 40), cfd.CavityFVM(61, 40)
    big = np.asfortranarray(rng.standard_normal((ref.ni + 3, ref.nj)))       # ld = ni + 3
    a = big[:ref.ni, :]
    assert a.strides == (8, 8 * (ref.ni + 3))
    ref.upload("v", np.asfortranarray(a))
    new.upload_noalloc("v", a)
    assert np.array_equal(new.download("v"), ref.download("v"))
    assert np.array_equal(new.download("v"), a)
    ref.close()
    new.close()


def test_noalloc_upload_allocates_no_device_memory(cfd):
    """The defining property: once the context exists, free device memory is the same before and after the call
    (the staging area was materialised by cfd_create, not by the upload)."""
    import torch
    g = cfd.CavityFVM(500, 300)
    a = np.asfortranarray(np.random.default_rng(3).standard_normal((g.ni, g.nj)))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    g.upload_noalloc("u", a)                                                  # first call of this context
    g.upload_noalloc("p", a)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    assert np.array_equal(g.download("u"), a)
    g.close()


def test_noalloc_upload_feeds_the_solver(cfd):
    g = cfd.CavityFVM(10, 10, order="sequential")
    h = cfd.CavityFVM(10, 10, order="sequential")
    g.init_lid_driven(reynolds=100.0)
    h.setgrid()
    for name in ("den", "vis", "u"):
        h.upload_noalloc(name, g.download(name))
    assert g.iterate(5) == h.iterate(5)
    assert np.array_equal(g.download("u"), h.download("u"))
    g.close()
    h.close()


def test_noalloc_upload_rejects_grids_beyond_its_limit(cfd):
    g = cfd.CavityFVM(2100, 2100)             # 2102*2102 nodes > 4 194 304
    a = np.zeros((g.ni, g.nj), order="F")
    with pytest.raises(cfd.CfdError, match="exceeds the built-in staging area"):
        g.upload_noalloc("u", a)
    g.upload("u", a)                          # the allocating entry point still takes it
    g.close()
