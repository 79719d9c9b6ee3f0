"""Ray-set file format: round trip incl. inf/NaN bit patterns, empty sets, malformed files."""
import numpy as np
import pytest
from b200rt import rayfile, scenes


def test_round_trip(tmp_path):
    r = scenes.random_rays([-1, -1, -1, 1, 1, 1], 1000, seed=3)
    r[5, 3] = np.float32(0.25)
    r[7, 4] = np.nan
    r[9, 4:7] = 0
    p = tmp_path / "r.rays"
    rayfile.save(p, r, any_hit=True)
    back, flags = rayfile.load(p)
    assert flags == rayfile.FLAG_ANY_HIT
    assert np.array_equal(back.view(np.uint32), r.view(np.uint32))


def test_empty_and_malformed(tmp_path):
    p = tmp_path / "e.rays"
    rayfile.save(p, np.zeros((0, 8), np.float32))
    back, flags = rayfile.load(p)
    assert back.shape == (0, 8) and flags == 0
    p.write_bytes(b"nonsense")
    with pytest.raises(ValueError):
        rayfile.load(p)
    rayfile.save(p, np.zeros((4, 8), np.float32))
    p.write_bytes(p.read_bytes()[:-5])
    with pytest.raises(ValueError):
        rayfile.load(p)
