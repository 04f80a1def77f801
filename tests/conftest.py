"""Shared fixtures.  `-m gpu` tests need a B200 and call the product through the C ABI;
everything else runs on CPU (oracle vs golden fixtures / vs the compiled reference when
oracle/_ref is present, host logic, ABI surface)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
FILM = (0.036, 0.036)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def cornell():
    from amber_b200 import cornel_box
    return cornel_box()


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import pyoracle
    pyoracle.build(ref=False)
    return pyoracle


@pytest.fixture(scope="session")
def oracle(cornell, oracle_lib):
    return oracle_lib.Oracle(cornell)


@pytest.fixture(scope="session")
def reference(cornell, oracle_lib):
    """The unmodified reference (oracle/_ref/libamber_ref.so); skipped where it was not built."""
    if not oracle_lib.reference_available():
        pytest.skip("oracle/_ref/libamber_ref.so not built (needs /root/reference)")
    return oracle_lib.Reference(cornell)


@pytest.fixture(scope="session")
def device():
    """One amber_cuda context on cuda:0.  Fails loudly if the library is missing."""
    from amber_b200.render import Device
    dev = Device(0)
    yield dev
    dev.close()


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def unit_vectors(rng, n):
    v = rng.normal(size=(n, 3)).astype(np.float32)
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)


def image_agreement(got, want, rtol=2e-3):
    """Fraction of pixels whose RGB agree within rtol of the pixel's own magnitude
    (plus a floor of 1e-6 of the image mean)."""
    floor = 1e-6 * max(float(np.abs(want).mean()), 1e-30)
    ok = np.abs(got - want) <= rtol * np.abs(want) + floor
    return float(ok.all(axis=-1).mean())


def diffuse_box(aperture_radius=0.01, extra=False):
    """A small low-variance scene for estimator-equivalence checks: closed grey box,
    one Phong wall, a large ceiling light, a diffuse sphere; thin lens inside the box.
    ``extra`` adds a glass sphere and a mirror sphere (every material kind on the path)."""
    from amber_b200 import FlatSceneBuilder
    b = FlatSceneBuilder()
    b.make_thin_lens([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0.9, 0, 0, 0, 1], 0.05, 1.5, aperture_radius, 5,
                     add_aperture_objects=True)

    def quad(m, p0, p1, p2, p3):
        b.add_triangle(p0, p1, p2, m)
        b.add_triangle(p2, p3, p0, m)

    grey = b.make_lambertian((.6, .6, .6))
    red = b.make_lambertian((.6, .2, .2))
    shiny = b.make_phong((.7, .7, .7), 16)
    light = b.make_diffuse_light((4.0, 4.0, 3.0))
    quad(grey, (-1, -1, 1), (1, -1, 1), (1, -1, -1), (-1, -1, -1))      # floor
    quad(grey, (-1, 1, 1), (-1, 1, -1), (1, 1, -1), (1, 1, 1))          # ceiling
    quad(red, (-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1))       # left
    quad(grey, (1, -1, 1), (1, 1, 1), (1, 1, -1), (1, -1, -1))          # right
    quad(shiny, (-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1))     # back
    quad(grey, (-1, -1, 1), (-1, 1, 1), (1, 1, 1), (1, -1, 1))          # front (behind camera)
    quad(light, (-0.5, 0.98, 0.5), (-0.5, 0.98, -0.5), (0.5, 0.98, -0.5), (0.5, 0.98, 0.5))  # faces down
    b.add_sphere((0.3, -0.6, -0.3), 0.4, b.make_lambertian((.3, .5, .7)))
    if extra:
        b.add_sphere((-0.45, -0.7, 0.1), 0.3, b.make_refraction(1.5))
        b.add_sphere((-0.3, 0.3, -0.6), 0.25, b.make_specular((.9, .9, .9)))
    return b.build()


def light_tracing_box():
    """The box with a wide aperture (radius 0.25): about one light path in a thousand lands on
    the lens inside the field of view (in the Cornell box: three in a hundred thousand), so
    light-tracing images hold hundreds of splats at test sizes."""
    return diffuse_box(aperture_radius=0.25, extra=True)


def sphere_light_box():
    """The wide-aperture box with an emitting SPHERE next to the ceiling light (and a brighter, smaller
    one): what LightSet::GenerateRay does with Sphere::SampleSurfacePoint
    (src/amber/scene/primitive_sphere.cc:115-122)."""
    from amber_b200 import FlatSceneBuilder
    b = FlatSceneBuilder()
    b.make_thin_lens([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0.9, 0, 0, 0, 1], 0.05, 1.5, 0.25, 5, add_aperture_objects=True)

    def quad(m, p0, p1, p2, p3):
        b.add_triangle(p0, p1, p2, m)
        b.add_triangle(p2, p3, p0, m)

    grey = b.make_lambertian((.6, .6, .6))
    quad(grey, (-1, -1, 1), (1, -1, 1), (1, -1, -1), (-1, -1, -1))
    quad(grey, (-1, 1, 1), (-1, 1, -1), (1, 1, -1), (1, 1, 1))
    quad(b.make_lambertian((.6, .2, .2)), (-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1))
    quad(grey, (1, -1, 1), (1, 1, 1), (1, 1, -1), (1, -1, -1))
    quad(b.make_phong((.7, .7, .7), 16), (-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1))
    quad(grey, (-1, -1, 1), (-1, 1, 1), (1, 1, 1), (1, -1, 1))
    quad(b.make_diffuse_light((2.0, 2.0, 1.5)), (-0.3, 0.98, 0.3), (-0.3, 0.98, -0.3), (0.3, 0.98, -0.3), (0.3, 0.98, 0.3))
    b.add_sphere((0.45, 0.1, -0.3), 0.2, b.make_diffuse_light((3.0, 2.0, 4.0)))
    b.add_sphere((-0.5, -0.5, 0.2), 0.08, b.make_diffuse_light((9.0, 9.0, 5.0)))
    b.add_sphere((-0.2, -0.7, -0.4), 0.3, b.make_lambertian((.3, .5, .7)))
    b.add_sphere((0.5, -0.75, 0.3), 0.25, b.make_refraction(1.5))
    return b.build()


def pinhole_box():
    """A closed box seen through MakePinholeLens (tilted, off-centre transform): row N1's second
    lens (src/amber/scene/lens_pinhole.cc)."""
    from amber_b200 import FlatSceneBuilder
    b = FlatSceneBuilder()
    b.make_pinhole_lens([0.8, 0, 0.6, 0.2, 0, 1, 0, 0.1, -0.6, 0, 0.8, 0.9, 0, 0, 0, 1], 0.04)

    def quad(m, p0, p1, p2, p3):
        b.add_triangle(p0, p1, p2, m)
        b.add_triangle(p2, p3, p0, m)

    grey = b.make_lambertian((.6, .6, .6))
    light = b.make_diffuse_light((4.0, 4.0, 3.0))
    quad(grey, (-1, -1, 1), (1, -1, 1), (1, -1, -1), (-1, -1, -1))
    quad(grey, (-1, 1, 1), (-1, 1, -1), (1, 1, -1), (1, 1, 1))
    quad(b.make_lambertian((.6, .2, .2)), (-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1))
    quad(grey, (1, -1, 1), (1, 1, 1), (1, 1, -1), (1, -1, -1))
    quad(b.make_phong((.7, .7, .7), 16), (-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1))
    quad(grey, (-1, -1, 1), (-1, 1, 1), (1, 1, 1), (1, -1, 1))
    quad(light, (-0.5, 0.98, 0.5), (-0.5, 0.98, -0.5), (0.5, 0.98, -0.5), (0.5, 0.98, 0.5))
    b.add_sphere((0.3, -0.6, -0.3), 0.4, b.make_refraction(1.5))
    return b.build()
