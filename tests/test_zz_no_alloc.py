"""amber_cuda_cast_static: the closest-hit entry point that allocates nothing (static device scratch), added
for callers that cannot call cudaMalloc.  Checked through the C ABI against
the existing amber_cuda_cast on the same rays -- the same kernels run, so the results must be bit-identical --
on two calls in a row (the second one smaller, so stale contents of the shared scratch would show), on the
constant-memory list (Cornell box) and on the wide BVH, and beyond the stated maximum it must refuse.

Collected last on purpose (a failure here hides nothing after it); it builds its own
context, so it passes alone in a fresh process and after the other tests alike."""
import numpy as np
import pytest

from conftest import unit_vectors

pytestmark = pytest.mark.gpu


def rays(rng, n):
    o = rng.uniform(-0.95, 0.95, size=(n, 3)).astype(np.float32)
    return o, unit_vectors(rng, n)


def cast_static(dev, o, d):
    """Inputs and outputs in caller-owned device memory (torch only holds the buffers)."""
    import torch
    n = o.shape[0]
    cuda = torch.device("cuda", dev.device)
    d_o, d_d = torch.from_numpy(o).to(cuda), torch.from_numpy(d).to(cuda)
    d_prim = torch.full((n,), -7, dtype=torch.int32, device=cuda)
    d_t = torch.full((n,), -7.0, dtype=torch.float32, device=cuda)
    torch.cuda.synchronize(cuda)
    dev.cast_static(d_o.data_ptr(), d_d.data_ptr(), n, d_prim.data_ptr(), d_t.data_ptr())
    return d_prim.cpu().numpy(), d_t.cpu().numpy()


@pytest.mark.parametrize("scene_name", ["cornell", "mesh"])
def test_static_scratch_cast_equals_the_allocating_cast(scene_name):
    from amber_b200 import Sensor, capi, cornel_box, cornel_mesh
    from amber_b200.capi import AmberCudaError
    from amber_b200.render import Device
    scene = cornel_box() if scene_name == "cornell" else cornel_mesh(subdivisions=3)  # 1280 + 18 triangles: wide BVH
    rng = np.random.default_rng(2026)
    with Device(0) as dev:
        dev.upload_scene(scene)
        dev.set_sensor(Sensor(32, 32))
        assert dev.stats()["accel"] == (0 if scene_name == "cornell" else 1)
        for n in (capi.CAST_STATIC_MAX, 4099):  # the maximum, then a smaller ragged call over the same scratch
            o, d = rays(rng, n)
            want_prim, want_t = dev.cast(o, d)
            prim, t = cast_static(dev, o, d)
            assert np.array_equal(prim, want_prim)
            assert np.array_equal(t, want_t, equal_nan=True)
            assert (want_prim >= 0).mean() > 0.5 and np.isnan(t[prim < 0]).all()
        # empty input is a no-op; one ray beyond the maximum is refused with a clear message and computes nothing
        dev.cast_static(0, 0, 0, 0, 0)
        with pytest.raises(AmberCudaError, match="exceed the static scratch"):
            dev.cast_static(1, 1, capi.CAST_STATIC_MAX + 1, 1, 1)
        o, d = rays(rng, 257)
        prim, t = cast_static(dev, o, d)  # the context still works after the refusal
        want_prim, want_t = dev.cast(o, d)
        assert np.array_equal(prim, want_prim) and np.array_equal(t, want_t, equal_nan=True)
