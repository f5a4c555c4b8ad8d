"""Widening row f3: the "opencv-object-detector-hog" FileStorage node (XML / YAML).  Host-only code of the C-ABI
library, pinned both ways against the UNMODIFIED reference (oracle/_ref/libcvref.so): files it writes load here,
files written here load there, and the text itself is the same."""
import os

import numpy as np
import pytest

from _oracle import Ref

pytestmark = pytest.mark.skipif(not Ref.available(), reason="oracle/_ref/libcvref.so not built")


@pytest.fixture(scope="module")
def ref():
    return Ref()


@pytest.mark.parametrize("ext", ["yml", "xml", "yaml"])
@pytest.mark.parametrize("which", [0, 1])
def test_reference_file_loads_here(ref, tmp_path, ext, which):
    from opencv_b200 import read_model
    det = ref.people_detector(which)                       # 3781 (64x128) / 1981 (48x96) floats
    g = dict(win=(64, 128) if which == 0 else (48, 96), block=(16, 16), stride=(8, 8), cell=(8, 8), nbins=9,
             win_sigma=-1.0, l2hys=0.2, gamma=1)
    path = str(tmp_path / ("people_%d.%s" % (which, ext)))
    ref.hog_save(path, det, g=g, nlevels=37)
    m = read_model(path)
    assert m["win_size"] == g["win"] and m["block_size"] == (16, 16) and m["block_stride"] == (8, 8)
    assert m["cell_size"] == (8, 8) and m["nbins"] == 9 and m["nlevels"] == 37 and m["gamma_correction"] is True
    assert m["win_sigma"] == 4.0 and m["l2hys_threshold"] == 0.2
    assert m["detector"].dtype == np.float32 and np.array_equal(m["detector"], det)        # bit-exact round trip
    assert np.array_equal(read_model(path, "people_%d" % which)["detector"], det)          # by object name


@pytest.mark.parametrize("ext", ["yml", "xml"])
def test_file_written_here_loads_in_reference_and_has_the_same_text(ref, tmp_path, ext):
    from opencv_b200 import write_model
    rng = np.random.RandomState(7)
    det = (rng.randn(3781) * np.exp(rng.randn(3781) * 3)).astype(np.float32)     # many magnitudes
    det[:6] = [0.0, 1.0, -2.0, 1e-30, -3.5e12, 0.5]
    mine, theirs = str(tmp_path / ("det." + ext)), str(tmp_path / "ref" / ("det." + ext))
    os.makedirs(os.path.dirname(theirs))
    write_model(mine, det, win_sigma=-1.0, l2hys_threshold=0.3, gamma_correction=False, nlevels=21)
    got = ref.hog_load(mine)
    assert got["geom"] == [64, 128, 16, 16, 8, 8, 8, 8, 9] and got["nlevels"] == 21 and got["gamma_correction"] is False
    assert got["win_sigma"] == 4.0 and got["l2hys_threshold"] == 0.3
    assert np.array_equal(got["detector"], det)
    g = dict(win=(64, 128), block=(16, 16), stride=(8, 8), cell=(8, 8), nbins=9, win_sigma=-1.0, l2hys=0.3, gamma=0)
    ref.hog_save(theirs, det, g=g, nlevels=21)
    assert open(mine).read().split() == open(theirs).read().split()               # same tokens in the same order
    assert open(mine).read() == open(theirs).read()                               # and the same layout


def test_no_detector_and_errors(ref, tmp_path):
    from opencv_b200 import HogError, read_model, write_model
    p = str(tmp_path / "3d-model.v2.yml")
    write_model(p, None, nlevels=5)
    assert "_3d-model_v2:" in open(p).read()               # FileStorage::getDefaultObjectName (persistence_cpp.cpp:192-229)
    assert read_model(p)["detector"].size == 0 and ref.hog_load(p)["detector"].size == 0
    with pytest.raises(HogError):
        read_model(str(tmp_path / "missing.yml"))
    with pytest.raises(HogError):
        read_model(p, "no_such_object")
    with pytest.raises(HogError):
        write_model(str(tmp_path / "x.json"), None)
    bad = str(tmp_path / "bad.yml")
    open(bad, "w").write(open(p).read().replace("signedGradient: 0", "signedGradient: 1"))
    with pytest.raises(HogError):
        read_model(bad)                                     # refused, not silently ignored
