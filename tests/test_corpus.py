"""Packed corpus + native minibatch loader (dg_corpus_*): host-only, runs without a GPU.
Parity: packing the reference's own Torch7 record files gives the same bytes as the reference-format reader record by
record, and a minibatch read from a corpus of all 4,398 fixture positions equals the fixture arrays."""
import os

import numpy as np
import pytest

import __graft_entry__ as entry

REC = os.path.join(entry.ROOT, "tests", "golden", "records")
GAME = "data/test/1993/2000-03-24b.sgf"
MOVES = (1, 2, 3, 60, 125)


@pytest.fixture(scope="module")
def data():
    entry.build()
    from deepgo_b200.host import data
    return data


def test_pack_torch7_records_and_read_back(data, golden, tmp_path):
    files = [os.path.join(REC, GAME, str(mv)) for mv in MOVES]
    path = data.Corpus.pack(files, tmp_path / "c.dgc", threads=3)
    assert os.path.getsize(path) == 64 + 3264 * len(files)
    c = data.Corpus(path, threads=2)
    assert c.count == len(files)
    got = c.read(np.arange(len(files)))
    sel = np.where(golden["split"] == 2)[0]
    for k, mv in enumerate(MOVES):
        i = sel[golden["move"][sel] == mv][0]
        assert np.array_equal(got["raw"][k], golden["raw"][i])
        assert (got["x"][k], got["y"][k], got["player"][k]) == (golden["x"][i], golden["y"][i], golden["player"][i])
        assert got["rank"][k] == golden["ranks"][i][golden["player"][i] - 1]       # data.ranks[data.move.player]
    # any order, repeats, no moves wanted
    got2 = c.read([4, 0, 0, 2], want_moves=False)
    assert got2["x"] is None and np.array_equal(got2["raw"], got["raw"][[4, 0, 0, 2]])


def test_whole_fixture_corpus(data, golden, tmp_path):
    path = data.write_corpus_from_arrays(tmp_path / "all.dgc", golden["raw"], golden["x"], golden["y"], golden["player"], golden["ranks"])
    c = data.Corpus(path, threads=4)
    assert c.count == len(golden["raw"]) == 4398
    rng = np.random.default_rng(0)
    for n in (1, 128, 2048):
        idx = rng.integers(0, c.count, n)
        b = c.read(idx)
        assert np.array_equal(b["raw"], golden["raw"][idx])
        assert np.array_equal(b["x"], golden["x"][idx]) and np.array_equal(b["y"], golden["y"][idx])
        assert np.array_equal(b["player"], golden["player"][idx])
        assert np.array_equal(b["rank"], golden["ranks"][idx, golden["player"][idx] - 1])
    mb = c.raw_minibatch(64, rng)
    assert mb["raw"].shape == (64, 9, 19, 19) and mb["player"].min() >= 1 and mb["player"].max() <= 2
    assert c.read(np.zeros(0, np.int64))["raw"].shape == (0, 9, 19, 19)          # empty batch


def test_errors(data, tmp_path):
    import deepgo_b200
    with pytest.raises(deepgo_b200.DeepGoError):
        data.Corpus(tmp_path / "missing.dgc")
    bad = tmp_path / "bad.dgc"
    bad.write_bytes(b"x" * 100)
    with pytest.raises(deepgo_b200.DeepGoError):
        data.Corpus(bad)
    with pytest.raises(deepgo_b200.DeepGoError):                                   # not a Torch7 record
        data.Corpus.pack([os.path.join(REC, "data", "test_game_counts.txt")], tmp_path / "x.dgc")
    files = [os.path.join(REC, GAME, "1")]
    c = data.Corpus(data.Corpus.pack(files, tmp_path / "one.dgc"))
    with pytest.raises(deepgo_b200.DeepGoError):
        c.read([1])                                                               # out of range
    with pytest.raises(deepgo_b200.DeepGoError):
        c.read([-1])
