"""The render queue, pixel windows, interleaved sample indices and the in-library film reduction (C ABI v2):
what the C++ / Python drop-in uses to keep G devices busy concurrently, as the reference keeps its worker
threads busy (src/amber/prelude/parallel.cc:29-40, include/amber/rendering/parallel.h:55-68)."""
import threading

import numpy as np
import pytest

from conftest import FILM, diffuse_box, image_agreement

pytestmark = pytest.mark.gpu


def close(a, b):
    return np.allclose(a, b, rtol=1e-4, atol=1e-7 * max(float(b.mean()), 1e-30))


@pytest.fixture
def cornell_device(device, cornell):
    if device.scene is not cornell:
        device.upload_scene(cornell)
    return device


def test_pixel_window_equals_a_crop_of_the_full_render(cornell_device):
    from amber_b200 import Sensor, capi
    from amber_b200.capi import AmberCudaError
    w, h = 96, 64
    cornell_device.set_sensor(Sensor(w, h))
    for estimator in (capi.ESTIMATOR_REFERENCE, capi.ESTIMATOR_NEE):
        cornell_device.film_clear()
        cornell_device.render(3, 8, seed=4, max_depth=16, estimator=estimator)
        full, _ = cornell_device.resolve()
        for first, count in ((16 * w, 16 * w), (5, 1), (w * h - 700, 700)):  # rows, one pixel, a ragged tail
            cornell_device.film_clear()
            before = cornell_device.stats()["paths"]
            cornell_device.render(3, 8, seed=4, max_depth=16, estimator=estimator, window=(first, count))
            part, done = cornell_device.resolve()
            assert done == 8 and cornell_device.stats()["paths"] - before == 8 * count
            flat_full, flat_part = full.reshape(-1, 3), part.reshape(-1, 3)
            assert close(flat_part[first:first + count], flat_full[first:first + count])
            assert not flat_part[:first].any() and not flat_part[first + count:].any()
    with pytest.raises(AmberCudaError, match="window"):
        cornell_device.render(0, 1, window=(w * h - 4, 8))
    with pytest.raises(AmberCudaError, match="window"):
        cornell_device.render(0, 1, estimator=capi.ESTIMATOR_LIGHT_TRACING, window=(0, 64))


def test_interleaved_sample_indices_cover_the_same_samples(cornell_device):
    """Device d of G renders the indices d, d + G, ...: two strided halves equal the contiguous whole."""
    from amber_b200 import Sensor, capi
    cornell_device.set_sensor(Sensor(80, 48))
    for estimator in (capi.ESTIMATOR_NEE, capi.ESTIMATOR_LIGHT_TRACING, capi.ESTIMATOR_BDPT):
        depth = 0 if estimator == capi.ESTIMATOR_BDPT else 12
        cornell_device.film_clear()
        cornell_device.render(0, 12, seed=6, max_depth=depth, estimator=estimator)
        whole, _ = cornell_device.resolve()
        cornell_device.film_clear()
        for d in range(3):
            cornell_device.render(d, 4, seed=6, max_depth=depth, estimator=estimator, sample_stride=3)
        strided, done = cornell_device.resolve()
        assert done == 12 and close(strided, whole), estimator


def test_queued_calls_run_as_one_wavefront(cornell_device):
    """amber_cuda_render_async: calls that continue the running one are absorbed into its wavefront (no pool
    drain in between) and give the image of one big call."""
    from amber_b200 import Sensor
    cornell_device.set_sensor(Sensor(1024, 1024))
    cornell_device.render(0, 16, seed=9, max_depth=32)
    whole, _ = cornell_device.resolve()
    chained0 = cornell_device.stats()["chained_renders"]
    cornell_device.film_clear()
    for k in range(8):
        cornell_device.render(2 * k, 2, seed=9, max_depth=32, wait=False)
    assert cornell_device.pending() >= 0
    cornell_device.wait()
    assert cornell_device.pending() == 0
    split, done = cornell_device.resolve()
    stats = cornell_device.stats()
    assert done == 16 and stats["paths"] == 16 * 1024 * 1024
    assert close(split, whole)
    assert stats["chained_renders"] - chained0 >= 1, "no queued call was absorbed into the running wavefront"
    # a call that does NOT continue the previous one (other seed) starts its own wavefront; order is kept
    cornell_device.film_clear()
    cornell_device.render(0, 2, seed=9, max_depth=32, wait=False)
    cornell_device.render(2, 2, seed=10, max_depth=32, wait=False)
    cornell_device.wait_pending(1)
    mixed, done = cornell_device.resolve()  # resolve implies wait
    assert done == 4 and np.isfinite(mixed).all()
    cornell_device.set_sensor(Sensor(64, 64))


def test_film_reduce_sums_the_films_of_several_contexts(cornell):
    """amber_cuda_film_reduce: contexts of one process (here on one device; on a multi-GPU box over peer
    memory) each render a share of the samples; the sum equals one context rendering all of them."""
    from amber_b200 import Sensor
    from amber_b200.capi import AmberCudaError
    from amber_b200.render import Device
    with Device(0) as a, Device(0) as b, Device(0) as c:
        for dev in (a, b, c):
            dev.upload_scene(cornell)
            dev.set_sensor(Sensor(64, 40))
        for d, dev in enumerate((a, b, c)):
            dev.render(d, 5, seed=3, max_depth=16, sample_stride=3, wait=False)   # all three at once
        a.film_reduce_from([b, c])
        image, done = a.resolve()
        assert done == 15
        b.film_clear()
        b.render(0, 15, seed=3, max_depth=16)
        whole, _ = b.resolve()
        assert close(image, whole)
        c.set_sensor(Sensor(32, 32))
        with pytest.raises(AmberCudaError, match="sensor"):
            a.film_reduce_from([c])


def test_two_threads_drive_two_contexts_on_one_device(cornell, oracle):
    """The scene of a context lives in the device's __constant__ banks while it renders; contexts that share
    a device take turns (per-device lock).  Two host threads, two scenes, many renders each: every image
    is the one its own scene gives."""
    from amber_b200 import Sensor
    from amber_b200.render import Device
    scenes = (cornell, diffuse_box(extra=True))
    want, errors = {}, []
    with Device(0) as a, Device(0) as b:
        devices = (a, b)
        for dev, scene in zip(devices, scenes):
            dev.upload_scene(scene)
            dev.set_sensor(Sensor(48, 48))
            dev.render(0, 4, seed=7, max_depth=8, estimator=0)
            want[dev] = dev.resolve()[0]

        def work(dev):
            try:
                for _ in range(20):
                    dev.film_clear()
                    dev.render(0, 4, seed=7, max_depth=8, estimator=0)
                    image, _ = dev.resolve()
                    if not close(image, want[dev]):
                        errors.append("image of the other context's scene")
                    o = np.zeros((64, 3), np.float32)
                    d = np.tile(np.array([[0, 0, -1]], np.float32), (64, 1))
                    dev.cast(o, d)
            except Exception as error:  # noqa: BLE001
                errors.append(repr(error))

        threads = [threading.Thread(target=work, args=(dev,)) for dev in devices]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not errors, errors[:3]
        reference, _ = oracle.render(48, 48, *FILM, 0, 4, 7, 8, estimator=0)
        assert image_agreement(want[a], reference) >= 0.99


def test_sample_count_travels_in_the_film_across_an_external_reduce(cornell_device):
    """The multi-process film reduce is ONE collective: the count rides in the film's trailing element and
    resolve reads it on the device.  Doubling the buffer stands in for a sum over two identical ranks."""
    import torch
    from amber_b200 import Sensor, capi
    from amber_b200.dist import film_tensor
    cornell_device.set_sensor(Sensor(40, 24))
    cornell_device.render(0, 70000, seed=2, max_depth=8)
    image, done = cornell_device.resolve()
    film = film_tensor(cornell_device)
    assert film.numel() == 4 * (40 * 24 + 1)
    assert film[-4:].tolist() == [float(done & 0xFFFF), float(done >> 16), 0.0, 0.0]
    film *= 2.0
    torch.cuda.synchronize()
    cornell_device.film_set_samples(capi.SAMPLES_IN_FILM)
    doubled, total = cornell_device.resolve()
    assert total == 2 * done and close(doubled, image)
    cornell_device.render(done, 2, seed=2, max_depth=8)     # more samples on top of a reduced film
    _, total = cornell_device.resolve()
    assert total == 2 * done + 2
    cornell_device.set_sensor(Sensor(64, 64))


def test_host_algorithm_on_several_devices(cornell):
    """CudaPathTracing(devices=[...]): the devices render concurrently and the image holds the same samples
    as one device's (skipped on a one-GPU box; the single-device queueing path runs either way)."""
    import time

    import torch
    from amber_b200 import Context, CudaPathTracing, Sensor
    sensor = Sensor(512, 512)
    single = CudaPathTracing(max_depth=16, seed=5, batch_spp=8, min_batch_paths=1)
    image1 = single.render(cornell, sensor, Context(1, 64))
    assert single.last_stats["paths"] == 64 * 512 * 512 and single.last_stats["chained_renders"] >= 1
    t0 = time.perf_counter()
    single.render(cornell, sensor, Context(1, 256))
    t_single = time.perf_counter() - t0
    single.close()
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU: multi-device dealing not exercised")
    n = min(4, torch.cuda.device_count())
    multi = CudaPathTracing(max_depth=16, seed=5, batch_spp=8, devices=list(range(n)), min_batch_paths=1)
    image_n = multi.render(cornell, sensor, Context(1, 64))
    assert multi.last_stats["paths"] == 64 * 512 * 512
    assert close(image_n, image1)
    t0 = time.perf_counter()
    multi.render(cornell, sensor, Context(1, 256))
    t_multi = time.perf_counter() - t0
    multi.close()
    print(f"256 spp at 512x512: 1 device {t_single * 1e3:.1f} ms, {n} devices {t_multi * 1e3:.1f} ms")
    assert t_multi < 0.75 * t_single
