"""Host-side mirror of the reference's interface for the path-tracing path.

The reference drives its path tracer as
``algorithm.Render(scene, sensor, context) -> Image<RGB>``
(``include/amber/rendering/algorithm.h:31-45``; called from
``src/amber/cli/application.cc:149-152``).  The same three nouns exist here with the
same meaning, so that the tests read like a user of the reference:

  * ``Sensor(width, height, film_width, film_height)`` -- ``rendering::Sensor``
    (``src/amber/rendering/sensor.cc:28-38``).
  * ``Context(n_threads, n_iterations)`` -- ``cli::Context``
    (``src/amber/cli/context.cc:34-67``): one successful ``iterate()`` is one sample
    per pixel over the whole frame; ``expire()`` ends the render early and the
    partial mean is returned.
  * ``CudaPathTracing(...).render(scene, sensor, context)`` -- the ``Algorithm<RGB>``
    whose body runs on the GPU through the C ABI (``include/amber_cuda.h``).

``Device`` is the thin object wrapper over one ``amber_cuda_ctx``.  There is no CPU
path: constructing a ``Device`` without the built library or without a GPU raises.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import capi
from .scene import FlatScene


class Sensor:
    """rendering::Sensor: pixel grid + film size in scene units."""

    def __init__(self, width: int, height: int, film_width: float = 0.036, film_height: float | None = None):
        self.width = int(width)
        self.height = int(height)
        self.film_width = float(film_width)
        # src/amber/cli/application.cc:89-94: 36 mm film, height scaled by aspect
        self.film_height = float(film_height) if film_height is not None else 0.036 * height / width

    def size(self) -> int:
        return self.width * self.height


class Context:
    """cli::Context: thread count, iteration budget, expiry."""

    def __init__(self, n_threads: int = 1, n_iterations: int = 0):
        self._lock = threading.Lock()
        self._n_threads = n_threads
        self._n_iterations = n_iterations
        self._started = 0
        self._expired = False

    def thread_count(self) -> int:
        return self._n_threads

    def iteration_count(self) -> int:
        with self._lock:
            return self._started

    def iterate(self) -> bool:
        with self._lock:
            if self._n_iterations > 0 and self._started >= self._n_iterations:
                return False
            if self._expired:
                return False
            self._started += 1
            return True

    def expire(self) -> None:
        with self._lock:
            self._expired = True


class Device:
    """One ``amber_cuda_ctx`` (one GPU, one stream)."""

    def __init__(self, device: int = 0, lib_path: str | None = None):
        self.lib = capi.load(lib_path)
        self.ctx = C.c_void_p()
        rc = self.lib.amber_cuda_create(C.byref(self.ctx), int(device))
        if rc != 0:
            msg = self.lib.amber_cuda_last_error(None)
            raise capi.AmberCudaError(f"amber_cuda_create failed ({rc}): {msg.decode() if msg else '?'}")
        self.device = int(device)
        self.sensor: Sensor | None = None
        self.scene: FlatScene | None = None

    def _check(self, rc: int) -> None:
        capi.check(self.lib, self.ctx, rc)

    def close(self) -> None:
        if self.ctx:
            self.lib.amber_cuda_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- scene / sensor -----------------------------------------------------------
    def upload_scene(self, scene: FlatScene) -> None:
        self._check(self.lib.amber_cuda_upload_scene(self.ctx, C.byref(scene.desc)))
        self.scene = scene

    def set_sensor(self, sensor: Sensor) -> None:
        self._check(self.lib.amber_cuda_set_sensor(self.ctx, sensor.width, sensor.height,
                                                   sensor.film_width, sensor.film_height))
        self.sensor = sensor

    def film_clear(self) -> None:
        self._check(self.lib.amber_cuda_film_clear(self.ctx))

    # -- rendering ----------------------------------------------------------------
    def render(self, sample_begin: int, sample_count: int, seed: int = 1, max_depth: int = 0,
               estimator: int = capi.ESTIMATOR_NEE, pool_paths: int = 0, window: tuple[int, int] = (0, 0),
               sample_stride: int = 1, wait: bool = True) -> None:
        """Add ``sample_count`` samples per pixel.  ``window`` = (first pixel, pixel count) restricts the
        render to a range of pixel indices; ``wait=False`` only queues the call on the context's worker
        thread (``amber_cuda_render_async``): consecutive queued calls run as one wavefront."""
        params = capi.RenderParams(sample_begin, sample_count, max_depth, seed, estimator, pool_paths,
                                   int(window[0]), int(window[1]), int(sample_stride), 0)
        call = self.lib.amber_cuda_render if wait else self.lib.amber_cuda_render_async
        self._check(call(self.ctx, C.byref(params)))

    def wait(self) -> None:
        """Block until every queued render is done; raises the queue's first failure."""
        self._check(self.lib.amber_cuda_wait(self.ctx))

    def pending(self) -> int:
        return int(self.lib.amber_cuda_pending(self.ctx))

    def wait_pending(self, max_pending: int) -> None:
        self._check(self.lib.amber_cuda_wait_pending(self.ctx, int(max_pending)))

    def resolve(self, out: np.ndarray | None = None) -> tuple[np.ndarray, int]:
        assert self.sensor is not None
        if out is None:
            out = np.empty((self.sensor.height, self.sensor.width, 3), dtype=np.float32)
        done = C.c_uint64()
        self._check(self.lib.amber_cuda_resolve(self.ctx, out.ctypes.data_as(C.c_void_p), C.byref(done)))
        return out, done.value

    def resolve_ldr(self, flip_x: bool = False) -> tuple[np.ndarray, int]:
        """Mean image through Filmic + Gamma, 8-bit RGB (H, W, 3): what the reference writes as
        ``<output>.png`` (``src/amber/cli/application.cc:98-106``); ``flip_x`` mirrors the
        columns as its ``ExportPNG`` does."""
        assert self.sensor is not None
        out = np.empty((self.sensor.height, self.sensor.width, 3), dtype=np.uint8)
        done = C.c_uint64()
        self._check(self.lib.amber_cuda_resolve_ldr(self.ctx, out.ctypes.data_as(C.c_void_p), int(flip_x), C.byref(done)))
        return out, done.value

    def film_device(self) -> tuple[int, int, int]:
        ptr, n, done = C.c_void_p(), C.c_uint64(), C.c_uint64()
        self._check(self.lib.amber_cuda_film_device(self.ctx, C.byref(ptr), C.byref(n), C.byref(done)))
        return ptr.value, n.value, done.value

    def film_set_samples(self, samples: int) -> None:
        self._check(self.lib.amber_cuda_film_set_samples(self.ctx, samples))

    def film_reduce_from(self, others: "list[Device]") -> None:
        """Sum the films of ``others`` (contexts of this process, any devices) into this one:
        ``amber_cuda_film_reduce`` (peer reads over NVLink)."""
        ctxs = (C.c_void_p * (1 + len(others)))(self.ctx, *[o.ctx for o in others])
        self._check(self.lib.amber_cuda_film_reduce(ctxs, len(ctxs)))

    # -- parity entry points ----------------------------------------------------------
    def cast(self, origins, directions) -> tuple[np.ndarray, np.ndarray]:
        o = np.ascontiguousarray(origins, dtype=np.float32)
        d = np.ascontiguousarray(directions, dtype=np.float32)
        n = o.shape[0]
        prim = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float32)
        self._check(self.lib.amber_cuda_cast(self.ctx, o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p),
                                             n, prim.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p)))
        return prim, t

    def cast_static(self, origins3_ptr: int, directions3_ptr: int, n: int, prim_ptr: int, t_ptr: int) -> None:
        """``amber_cuda_cast_static``: closest hits with device-resident inputs / outputs and no allocation
        inside the library (static scratch, at most ``capi.CAST_STATIC_MAX`` rays per call)."""
        self._check(self.lib.amber_cuda_cast_static(self.ctx, origins3_ptr, directions3_ptr, n, prim_ptr, t_ptr))

    def cast_device(self, origin4_ptr: int, direction4_ptr: int, n: int, hit4_ptr: int) -> float:
        ms = C.c_float()
        self._check(self.lib.amber_cuda_cast_device(self.ctx, origin4_ptr, direction4_ptr, n, hit4_ptr, C.byref(ms)))
        return ms.value

    def sample_light(self, material: int, normals, dirs_out, xi8):
        nrm = np.ascontiguousarray(normals, dtype=np.float32)
        wo = np.ascontiguousarray(dirs_out, dtype=np.float32)
        xi = np.ascontiguousarray(xi8, dtype=np.float32)
        n = nrm.shape[0]
        wi = np.empty((n, 3), dtype=np.float32)
        w = np.empty((n, 3), dtype=np.float32)
        draws = np.empty(n, dtype=np.int32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.amber_cuda_sample_light(self.ctx, material, n, p(nrm), p(wo), p(xi), p(wi), p(w), p(draws)))
        return wi, w, draws

    def sample_importance(self, material: int, normals, dirs_out, xi8):
        """Material::SampleImportance (light tracing) with scripted uniforms."""
        nrm = np.ascontiguousarray(normals, dtype=np.float32)
        wo = np.ascontiguousarray(dirs_out, dtype=np.float32)
        xi = np.ascontiguousarray(xi8, dtype=np.float32)
        n = nrm.shape[0]
        wi = np.empty((n, 3), dtype=np.float32)
        w = np.empty((n, 3), dtype=np.float32)
        draws = np.empty(n, dtype=np.int32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.amber_cuda_sample_importance(self.ctx, material, n, p(nrm), p(wo), p(xi), p(wi), p(w), p(draws)))
        return wi, w, draws

    def light_rays(self, xi5):
        """Scene::GenerateLightRay with scripted uniforms: origins, directions, rgb weights."""
        xi = np.ascontiguousarray(xi5, dtype=np.float32)
        n = xi.shape[0]
        o = np.empty((n, 3), dtype=np.float32)
        d = np.empty((n, 3), dtype=np.float32)
        w = np.empty((n, 3), dtype=np.float32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.amber_cuda_light_rays(self.ctx, n, p(xi), p(o), p(d), p(w)))
        return o, d, w

    def lens_response(self, positions, dirs_out):
        """Scene::Response: pixel index (x + y*W, -1 = none) and response value per arriving ray."""
        pos = np.ascontiguousarray(positions, dtype=np.float32)
        wo = np.ascontiguousarray(dirs_out, dtype=np.float32)
        n = pos.shape[0]
        pixel = np.empty(n, dtype=np.int32)
        value = np.empty(n, dtype=np.float32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.amber_cuda_lens_response(self.ctx, n, p(pos), p(wo), p(pixel), p(value)))
        return pixel, value

    def normalize(self, vectors, denominators=None):
        """prelude::Normalize on the device (the exact normalisation of vec.cuh); with `denominators`
        the IEEE quotients vectors[i] / denominators[i] instead.  For the parity tests."""
        v = np.ascontiguousarray(vectors, dtype=np.float32)
        out = np.empty_like(v)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        den = None if denominators is None else np.ascontiguousarray(denominators, dtype=np.float32)
        self._check(self.lib.amber_cuda_normalize(self.ctx, v.shape[0], p(v), None if den is None else p(den), p(out)))
        return out

    def eye_rays(self, px, py, xi5):
        px_ = np.ascontiguousarray(px, dtype=np.uint32)
        py_ = np.ascontiguousarray(py, dtype=np.uint32)
        xi = np.ascontiguousarray(xi5, dtype=np.float32)
        n = px_.shape[0]
        o = np.empty((n, 3), dtype=np.float32)
        d = np.empty((n, 3), dtype=np.float32)
        w = np.empty(n, dtype=np.float32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.amber_cuda_eye_rays(self.ctx, n, p(px_), p(py_), p(xi), p(o), p(d), p(w)))
        return o, d, w

    def stats(self) -> dict:
        s = capi.Stats()
        self._check(self.lib.amber_cuda_get_stats(self.ctx, C.byref(s)))
        return {name: getattr(s, name) for name, _ in capi.Stats._fields_}

    def set_profiling(self, on: bool) -> None:
        self._check(self.lib.amber_cuda_set_profiling(self.ctx, int(on)))


class CudaPathTracing:
    """rendering::Algorithm<RGB> for ``--algorithm pt`` on the GPU.

    ``render(scene, sensor, context)`` consumes ``context.iterate()`` in batches of
    ``batch_spp`` samples (so ``--spp``, expiry and the iteration counter keep their
    reference meaning) and returns the mean image, shape (H, W, 3), float32, row-major
    ``x + y*W`` like ``Image<RGB>`` (``include/amber/prelude/image.h:117-124``).

    Batches are queued (``amber_cuda_render_async``), at most ``depth`` ahead of the device, so
    consecutive batches run as ONE wavefront and the path pool is never drained in between; with
    several ``devices`` batch b goes to device b mod G, which renders the interleaved sample
    indices {d, d + G, ...}, all devices work concurrently -- as the reference's threads do
    (``include/amber/rendering/parallel.h:55-68``) -- and the films are summed over peer memory
    (``amber_cuda_film_reduce``).  Contexts, scene and pool stay alive between ``render`` calls.
    """

    def __init__(self, estimator: str = "nee", max_depth: int = 0, seed: int = 1, device: int = 0,
                 batch_spp: int = 64, pool_paths: int = 0, devices: "list[int] | None" = None, depth: int = 3,
                 min_batch_paths: int = 1 << 24):
        self.estimator = {"reference": capi.ESTIMATOR_REFERENCE, "nee": capi.ESTIMATOR_NEE,
                          "lt": capi.ESTIMATOR_LIGHT_TRACING, "bdpt": capi.ESTIMATOR_BDPT}[estimator]
        self.max_depth = max_depth
        self.seed = seed
        self.device_indices = list(devices) if devices else [device]
        self.batch_spp = max(1, batch_spp)
        self.pool_paths = pool_paths
        self.depth = max(1, depth)
        self.min_batch_paths = min_batch_paths  # small frames: a queued call holds at least this many camera paths
        self.last_stats: dict | None = None
        self._devices: list[Device] = []
        self._scene = None

    def close(self) -> None:
        for dev in self._devices:
            dev.close()
        self._devices = []
        self._scene = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _bind(self, scene: FlatScene, sensor: Sensor) -> "list[Device]":
        if not self._devices:
            self._devices = [Device(i) for i in self.device_indices]
        for dev in self._devices:
            if self._scene is not scene:
                dev.upload_scene(scene)
            dev.set_sensor(sensor)  # clears the film
        self._scene = scene
        return self._devices

    def render(self, scene: FlatScene, sensor: Sensor, context: Context) -> np.ndarray:
        if not isinstance(scene, FlatScene):
            # the reference adapter throws std::invalid_argument for a scene it cannot flatten
            raise TypeError("CudaPathTracing needs a FlatScene (no CPU fallback)")
        devices = self._bind(scene, sensor)
        n = len(devices)
        done = [0] * n  # samples per pixel handed to each device
        batch = 0
        batch_spp = min(1 << 20, max(self.batch_spp, -(-self.min_batch_paths // sensor.size())))
        while True:
            count = 0
            while count < batch_spp and context.iterate():
                count += 1
            if count == 0:
                break
            d = batch % n
            devices[d].wait_pending(self.depth - 1)
            devices[d].render(d + n * done[d], count, seed=self.seed, max_depth=self.max_depth,
                              estimator=self.estimator, pool_paths=self.pool_paths, sample_stride=n, wait=False)
            done[d] += count
            batch += 1
        for dev in devices:
            dev.wait()
        if n > 1:
            devices[0].film_reduce_from(devices[1:])
        image, _ = devices[0].resolve()
        self.last_stats = devices[0].stats()
        if n > 1:
            for key in ("paths", "closest_rays", "shadow_rays", "truncated_paths"):
                self.last_stats[key] = sum(dev.stats()[key] for dev in devices)
        return image


class CudaLightTracing(CudaPathTracing):
    """rendering::Algorithm<RGB> for ``--algorithm lt`` (row N2): the light tracer of
    ``src/amber/rendering/algorithm_lt.cc:82-163`` on the same wavefront kernels.  One
    ``context.iterate()`` is W*H light paths, as ``Thread::operator()`` traces them."""

    def __init__(self, max_depth: int = 0, seed: int = 1, device: int = 0, batch_spp: int = 64,
                 pool_paths: int = 0, devices: "list[int] | None" = None):
        super().__init__("lt", max_depth, seed, device, batch_spp, pool_paths, devices)


class CudaBidirectionalPathTracing(CudaPathTracing):
    """rendering::Algorithm<RGB> for ``--algorithm bdpt`` (row N3):
    ``src/amber/rendering/algorithm_bdpt.cc:82-148`` over
    ``include/amber/rendering/bidirectional_path_sampling.h``.  ``max_vertices`` caps the events
    per subpath (<= 16; the reference has no cap)."""

    def __init__(self, max_vertices: int = 0, seed: int = 1, device: int = 0, batch_spp: int = 64,
                 devices: "list[int] | None" = None):
        super().__init__("bdpt", max_vertices, seed, device, batch_spp, 0, devices)
