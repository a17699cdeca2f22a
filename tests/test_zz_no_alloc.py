"""rtd_add_device_to_host_noalloc (include/rtdose.h): the multi-GPU tail for callers that must not allocate. No counterpart in the
reference. Collected last on purpose.

Called through the C ABI and compared with the existing entry point rtd_add_device_to_host on the same inputs, two calls in a row (the
second call finds the scratch as the first one left it): both compute dst[i] += src[i] where src[i] != 0, one fp32 add per voxel, so
the grids must agree bit for bit and the counts exactly. The file needs nothing from the other tests (own context, own data)."""
import ctypes as C

import numpy as np
import pytest

from raytracedicom_b200 import engine

pytestmark = pytest.mark.gpu

MAX_VOXELS = 1 << 28        # RTD_NOALLOC_MAX_VOXELS


def _inputs(torch, n, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    src = torch.rand(n, generator=g, dtype=torch.float32)
    src[torch.rand(n, generator=g) < 0.7] = 0.0                 # a dose grid is mostly empty
    base = torch.rand(n, generator=g, dtype=torch.float32)
    return src.cuda(), base


@pytest.mark.parametrize("n", [1, 65536 * 3 + 17, 5_000_001])
def test_noalloc_tail_matches_the_allocating_entry_point_twice_in_a_row(n):
    import torch
    with engine.DoseEngine(0) as eng:
        for call in range(2):
            src, base = _inputs(torch, n, seed=100 + call)
            want = base.clone().pin_memory()
            got = base.clone().pin_memory()
            n_want = eng.add_device_to_host(src.data_ptr(), want.data_ptr(), n)
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info()[0]
            n_got = eng.add_device_to_host_noalloc(src.data_ptr(), got.data_ptr(), n)
            free1 = torch.cuda.mem_get_info()[0]
            assert n_got == n_want == int((src != 0).sum().item()), call
            assert np.array_equal(got.numpy().view(np.uint32), want.numpy().view(np.uint32)), call
            if call == 1:                                        # (the first launch of a kernel may load its code: checked on the second)
                assert free1 == free0, "the call changed the amount of free device memory"
        # a device destination is accepted too
        src, base = _inputs(torch, n, seed=7)
        d = base.cuda()
        m = eng.add_device_to_host_noalloc(src.data_ptr(), d.data_ptr(), n)
        assert m == int((src != 0).sum().item())
        assert np.array_equal(d.cpu().numpy().view(np.uint32), (base + src.cpu()).numpy().view(np.uint32))


def test_noalloc_tail_states_its_limits():
    import torch
    with engine.DoseEngine(0) as eng:
        src = torch.zeros(16, device="cuda")
        pinned = torch.zeros(16).pin_memory()
        with pytest.raises(engine.RtdError) as e:               # beyond the fixed scratch: refused before anything is touched
            eng.add_device_to_host_noalloc(src.data_ptr(), pinned.data_ptr(), MAX_VOXELS + 1)
        assert "2^28" in str(e.value) and "RTD_ERR_INVALID" in str(e.value)
        pageable = np.zeros(16, np.float32)
        with pytest.raises(engine.RtdError) as e:               # a pageable grid would need a staging buffer
            eng.add_device_to_host_noalloc(src.data_ptr(), pageable.ctypes.data, 16)
        assert "page-locked" in str(e.value)
        m = C.c_int64(-1)
        assert eng.lib.rtd_add_device_to_host_noalloc(eng._h, C.c_void_p(src.data_ptr()), C.c_void_p(pinned.data_ptr()), 0, C.byref(m)) == 0 and m.value == 0
