"""The port oracle (oracle/raycast_oracle.cpp) against the committed golden vectors, which were
produced by the reference's own compiled functions (tests/golden/make_golden.py).  Bit-exact."""
import json
import os

import numpy as np
import pytest

from parity import GOLDEN, default_origins, kat_names, load_npz_scene

DIR = np.array([0, 0, -1, -1], np.float32)


def test_scene1_full_frame_known_answer(port):
    info = json.load(open(os.path.join(GOLDEN, "scene1_640x480.json")))
    z = np.load(os.path.join(GOLDEN, "scene1.npz"))
    sc = load_npz_scene(z)
    rgba, pid, tt = port.render(sc, 640, 480, want_t=True)
    assert "%016x" % port.fnv1a64_words(rgba) == info["fnv1a64_words"] == "57116a151211b387"
    assert "%016x" % port.fnv1a64(rgba) == info["fnv1a64_bytes"]
    assert int((rgba[..., :3] != 0).any(-1).sum()) == info["pixels_with_nonzero_rgb"] == 38285
    assert int((rgba[..., :3] > 255).any(-1).sum()) == info["pixels_with_channel_gt_255"] == 6
    assert int(rgba.astype(np.int64).sum()) == info["sum_of_all_ints"]
    assert np.array_equal(rgba[::4, ::4], z["sub4"])
    hit = pid >= 0
    assert int(hit.sum()) == 38285 and int((tt[hit] < 0).sum()) == 25  # SURVEY.md s4 probe values
    assert abs(float(tt[hit].min()) - (-1.776)) < 1e-3 and abs(float(tt[hit].max()) - 129.97) < 1e-2
    assert (rgba[..., 3] == 255).all()


@pytest.mark.parametrize("name", ["scene3_seed7_window", "scene2_seed3_window"])
def test_reference_scene_windows(port, name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    sc = load_npz_scene(z)
    h, w = z["rgba"].shape[:2]
    rgba, _, _ = port.render(sc, w, h, ray_origins=default_origins(w, h, int(z["x0"]), int(z["y0"])), ray_dir=DIR)
    assert np.array_equal(rgba, z["rgba"])


@pytest.mark.parametrize("name", kat_names())
def test_known_answer_cases(port, name):
    z = np.load(os.path.join(GOLDEN, "kats.npz"))
    sc = load_npz_scene(z, name + "/")
    rgba, pid, tt = port.render(sc, 64, 48, want_t=True)
    assert np.array_equal(rgba, z[name + "/rgba"])
    # what each case is there to show
    if name == "axis_aligned_cube_pixel_exact_edges":
        assert (pid[10:31, 20:41] == 0).all() and (pid >= 0).sum() == 21 * 21 and tt[20, 30] == 60.0
    if name == "sphere_origin_inside_negative_t0":
        assert tt[24, 32] == -15.0 and rgba[24, 32, 1] == 276
    if name == "sphere_behind_plane_tca_negative" or name == "empty_scene":
        assert (pid == -1).all() and (rgba == [0, 0, 0, 255]).all()
    if name == "sphere_t0_exactly_zero_is_a_miss":
        assert pid[24, 32] == 1 and pid[24, 33] == 0  # centre pixel falls through to the far sphere
    if name == "cube_behind_plane_negative_t_wins":
        assert pid[20, 20] == 0 and tt[20, 20] == -48.0
    if name == "equal_t_cube_before_sphere":
        assert pid[24, 36] == 0 and tt[24, 36] == 50.0
    if name == "equal_t_lower_index_wins":
        assert pid[24, 30] == 2 and pid[10, 50] == 0
    if name == "t_beyond_180_wraps":
        assert tt[24, 20] == 242.0 and rgba[24, 20, 1] < 0


def test_single_function_vectors(port):
    z = np.load(os.path.join(GOLDEN, "function_vectors.npz"))
    for row, want in zip(z["sph_in"], z["sph_out"]):
        got = port.intersect_sphere(row[0:4], row[4:8], row[8], row[9:13])
        assert np.float32(got) == np.float32(want)
    n_hit = 0
    for row, want in zip(z["tri_in"], z["tri_out"]):
        hit, t = port.intersect_tri(row[0:3], row[3:6], row[6:9], row[9:12], row[12:15])
        assert hit == int(want[0])
        if hit:
            n_hit += 1
            assert t == want[1]
    assert n_hit > 0
