"""The command line's host side (no GPU): the reference's arguments (estimate_vel_vid.py:38-53) parse to the same meaning, the
defaults are the reference's pipeline, and the -c cache falls back to images.npy where h5py is missing."""
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_reference_arguments_and_defaults():
    cli = _load("estimate_vel_vid")
    a, whole = cli.parse(["10", "250", "-p", "/tmp/rec.d"])
    assert a.fileIndex == [10, 250] and not whole and str(a.srcDirectory) == "/tmp/rec.d"
    assert (a.continuation, a.h264) == ("False", "False")                    # the reference's string flags
    assert not (a.no_preprocess or a.no_clahe or a.no_denoise)               # CLAHE -> NLM -> pair loop, like the reference's main()
    assert (a.block, a.search_range, a.range_y, a.min_range, a.units) == (16, 64, 0, 6, "reference")
    a, whole = cli.parse(["-f", "-c", "-h264", "--no-preprocess", "--range", "32", "--range-y", "8"])
    assert whole and (a.continuation, a.h264) == ("True", "True") and a.no_preprocess and (a.search_range, a.range_y) == (32, 8)
    with pytest.raises(SystemExit):
        cli.parse(["10"])                                                    # FIRST LAST or -f
    raw = _load("estimate_vel")
    assert hasattr(raw, "main")


def test_cache_round_trip_without_h5py(tmp_path):
    cli = _load("estimate_vel_vid")
    frames = np.random.default_rng(0).integers(0, 256, (5, 4, 6), dtype=np.uint8)
    h5, npy = str(tmp_path / "images.h5"), str(tmp_path / "images.npy")
    assert cli.load_cache(h5, npy) is None
    cli.save_cache(frames, h5, npy)
    try:
        import h5py  # noqa: F401
        assert os.path.exists(h5)
    except ImportError:
        assert os.path.exists(npy) and not os.path.exists(h5)
    assert np.array_equal(cli.load_cache(h5, npy), frames)


def test_ranks_outside_torchrun(monkeypatch):
    cli = _load("estimate_vel_vid")
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    assert cli.ranks() == (0, 1)
