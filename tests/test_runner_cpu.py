"""CPU: the host stage of the batched caller loop (runner.py; run_mars_depth.py:90-117): decode + LANCZOS resize, source
detection, shape/source bucketing — against the reference's own lines restated with PIL."""
import os

import numpy as np
import pytest
from PIL import Image

from monodepth_estimation_b200 import runner


def _write_images(tmp, specs):
    paths = []
    rng = np.random.Generator(np.random.PCG64(11))
    for name, (w, h) in specs:
        arr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        p = os.path.join(tmp, name)
        Image.fromarray(arr).save(p)
        paths.append(p)
    return paths


def _reference_load(path, max_size):
    """run_mars_depth.py:94-100, verbatim steps."""
    img = Image.open(path).convert("RGB")
    if max(img.size) > max_size:
        scale = max_size / max(img.size)
        new_size = (int(img.size[0] * scale), int(img.size[1] * scale))
        img = img.resize(new_size, Image.LANCZOS)
    return np.array(img)


def test_source_detection_matches_reference_rules():
    assert runner.detect_source("/data/Rover_0012.png") == "rover"
    assert runner.detect_source("/data/ingenuity/frame.png") == "ingenuity"
    assert runner.detect_source("/data/AERIAL_7.jpg") == "ingenuity"
    assert runner.detect_source("/x/orbital_3.png") == "satellite"
    assert runner.detect_source("/x/satellite-rover.png") == "rover"  # the reference checks "rover" first (:107)
    assert runner.detect_source("/x/unknown.png") == "rover"
    assert runner.detect_source("/x/rover.png", "satellite") == "satellite"


def test_decode_resize_and_bucketing(tmp_path):
    specs = [("rover_a.png", (64, 48)), ("rover_b.png", (64, 48)), ("aerial_c.png", (64, 48)), ("rover_d.png", (300, 200)), ("rover_e.png", (64, 48)),
             ("rover_f.png", (150, 100)), ("rover_g.jpg", (64, 48))]
    paths = _write_images(str(tmp_path), specs)
    for p in paths:
        assert np.array_equal(runner.load_image_u8(p, 128), _reference_load(p, 128))
    assert runner.load_image_u8(paths[3], 128).shape == (85, 128, 3)  # int(200 * 128/300) = 85
    got = list(runner.iter_batches(paths, batch_size=2, max_size=128, workers=3))
    seen = sorted(i for idx, _, _ in got for i in idx)
    assert seen == list(range(len(paths)))  # every image exactly once
    for idx, batch, src in got:
        assert batch.dtype == np.uint8 and batch.ndim == 4 and batch.shape[0] == len(idx) <= 2
        for j, i in enumerate(idx):
            assert np.array_equal(batch[j], _reference_load(paths[i], 128))
            assert runner.detect_source(paths[i]) == src
    # full buckets come out as soon as they fill (arrival order), partial ones at the end
    assert [idx for idx, _, _ in got] == [[0, 1], [3, 5], [4, 6], [2]]  # 300x200 and 150x100 both become 128x85
    with pytest.raises(ValueError):
        list(runner.iter_batches(paths, batch_size=0))
