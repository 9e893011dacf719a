"""Port oracle vs oracle/_ref (the reference's own function bodies, compiled here).  Skipped where
the reference sources are absent (the GPU box); the committed golden vectors cover that case."""
import numpy as np

from parity import default_origins, to_oracle_scene

DIR = np.array([0, 0, -1, -1], np.float32)


def test_default_ray_dir_is_0_0_m1_m1(ref):
    assert ref.default_ray_dir().tolist() == [0.0, 0.0, -1.0, -1.0]  # MainState.cpp:37-39


def test_reference_scenes_bit_exact(ref, port):
    for which, seed, (w, h) in ((1, 1, (320, 240)), (2, 11, (160, 120)), (3, 5, (96, 64))):
        sc = ref.build_scene(which, seed)
        want = ref.render(default_origins(w, h), DIR, (h, w, 4))
        got, _, _ = port.render(sc, w, h)
        assert np.array_equal(got, want), (which, seed)


def test_faithful_variant_same_pixels(ref, port):
    sc = ref.build_scene(3, 9)
    a, _, _ = port.render(sc, 48, 32)
    b, _, _ = port.render(sc, 48, 32, faithful=True)
    assert np.array_equal(a, b)


def test_synthetic_scene_and_rotated_camera(ref, port, built_libs):
    from opencl_ray_tracer_b200 import host_scene as hs
    s = hs.random_scene(256, 128, 24, 24, seed=42)
    # feed the reference the same flattened scene through its own Cube class is impossible
    # (vertices are inputs here), so compare through explicit origins on both sides instead:
    ref.clear_scene()
    for o, r, c in zip(s.sph_o, s.sph_r, s.sph_c):
        ref.add_sphere(o, r, c)
    sc = to_oracle_scene(s)
    sc_spheres_only = type(sc)(s.sph_o, s.sph_r, s.sph_c, np.zeros((0, 36, 4), np.float32), np.zeros((0, 4), np.float32))
    cam = hs.orbit_camera(256, 128, 7)
    origins = port.pixel_origins(96, 64, *cam[:3])
    want = ref.render(origins, cam[3], (64, 96, 4))
    got, _, _ = port.render(sc_spheres_only, 96, 64, o0=cam[0], du=cam[1], dv=cam[2], ray_dir=cam[3])
    assert np.array_equal(got, want)
