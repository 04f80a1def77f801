"""The C++ drop-in (include/amber_cuda/adapter.hpp, amber::cuda::FlatSceneBuilder /
MakeRGBCudaPathTracing) linked against the compiled reference: oracle/_ref/adapter_demo
(built by `make -C oracle adapter_demo`, which needs /root/reference; the binary travels to
the GPU box with oracle/_ref/)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

DEMO = os.path.join(ROOT, "oracle", "_ref", "adapter_demo")
needs_demo = pytest.mark.skipif(not os.path.exists(DEMO), reason="oracle/_ref/adapter_demo not built (needs /root/reference)")


@needs_demo
def test_cpp_builder_records_the_same_flat_scene_as_the_python_builder(tmp_path, cornell):
    """Same factory calls -> byte-identical geometry, materials and lens on both host sides
    (the aperture blades are read back from the reference's own lens object)."""
    from amber_b200 import capi
    out = tmp_path / "flat.bin"
    subprocess.check_call([DEMO, "--dump", str(out)])
    record = np.dtype([("kind", "<i4"), ("data", "<f4", (9,)), ("mkind", "<i4"), ("rgb", "<f4", (3,)), ("param", "<f4")])
    raw = out.read_bytes()
    n = cornell.desc.n_primitives
    got = np.frombuffer(raw[: n * record.itemsize], dtype=record)
    assert len(raw) == n * record.itemsize + ctypes.sizeof(capi.Lens)
    # the aperture blades (Eye material) come first on both sides; the reference's lens hands them
    # out in its own order, so they are compared as a set, everything else position by position
    def eye_rows(kinds, rows):
        return sorted(tuple(r) for k, r in zip(kinds, rows) if k == capi.MAT_EYE)
    want_kinds = [cornell.materials[cornell.primitives[i].material].kind for i in range(n)]
    want_rows = [tuple(np.array(cornell.primitives[i].data[:], dtype=np.float32)) for i in range(n)]
    assert eye_rows(got["mkind"], [tuple(r) for r in got["data"]]) == eye_rows(want_kinds, want_rows)
    assert list(got["mkind"]) == want_kinds
    for i in range(n):
        p = cornell.primitives[i]
        m = cornell.materials[p.material]
        if m.kind == capi.MAT_EYE:
            continue
        assert got["kind"][i] == p.kind
        width = 9 if p.kind == capi.PRIM_TRIANGLE else 4
        assert np.array_equal(got["data"][i][:width], np.array(p.data[:width], dtype=np.float32)), i
        assert got["mkind"][i] == m.kind
        if m.kind not in (capi.MAT_REFRACTION, capi.MAT_EYE):
            assert np.array_equal(got["rgb"][i], np.array(m.rgb[:], dtype=np.float32)), i
        assert got["param"][i] == np.float32(m.param)
    lens = capi.Lens.from_buffer_copy(raw[n * record.itemsize:])
    assert bytes(lens) == bytes(cornell.desc.lens)


@needs_demo
def test_cpp_adapter_fails_loudly_without_a_gpu():
    """No CPU fallback: without a CUDA device the GPU algorithm throws (exit code 3)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    rc = subprocess.run([DEMO, "32", "2", "2"], capture_output=True, text=True)
    assert rc.returncode == 3 and "no CUDA device" in rc.stdout


@pytest.mark.gpu
@needs_demo
def test_cpp_adapter_renders_through_the_reference_interface():
    """rendering::Algorithm<RGB>::Render with the reference's CPU path tracer and with the GPU one
    on the same amber::cuda::Scene: block means agree, a non-flat scene is refused."""
    rc = subprocess.run([DEMO, "128", "1024", "16"], capture_output=True, text=True, timeout=600)
    print(rc.stdout)
    assert rc.returncode == 0, rc.stdout + rc.stderr
    assert "adapter_demo: OK" in rc.stdout and "refused as expected" in rc.stdout


@pytest.mark.gpu
@needs_demo
def test_cpp_adapter_queues_batches_as_one_wavefront_and_uses_every_device():
    """Algorithm<RGB>::Render on the GPU path tracer alone: batches of Iterate() are queued and chained (no
    pool drain between them), contexts persist across Render() calls, and with more than one device the
    image holds the same samples as one device's (checked inside the demo)."""
    import torch
    devices = min(2, torch.cuda.device_count())
    rc = subprocess.run([DEMO, "--throughput", "512", "256", str(devices), "16", "2"], capture_output=True, text=True, timeout=600)
    print(rc.stdout)
    assert rc.returncode == 0, rc.stdout + rc.stderr
    assert "adapter_demo: OK" in rc.stdout
    chained = [int(line.split('"chained_renders": ')[1].rstrip("}\n")) for line in rc.stdout.splitlines() if "chained_renders" in line]
    assert chained and max(chained) >= 1
