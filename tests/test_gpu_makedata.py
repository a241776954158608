"""-m gpu: SGF text -> records on the device path (dg_sgf_parse + lockstep dg_update_boards + one dg_summarize_boards): the records of
the fixture games, rebuilt from SGF text synthesised out of their own moves, must equal the reference's records bit for bit."""
import numpy as np
import pytest

from tests import gpu_util as U
from tests.test_makedata import sgf_of

pytestmark = pytest.mark.gpu


def test_fixture_games_rebuilt_on_the_gpu(golden, tmp_path):
    from deepgo_b200.host import makedata, data
    games = list(range(len(golden["games"])))
    texts, sels = zip(*[sgf_of(golden, gi) for gi in games])
    rec = makedata.transcribe(list(texts), U.ctx())
    order = np.concatenate(sels)
    assert len(rec["raw"]) == len(order) == 4398
    assert np.array_equal(rec["raw"], golden["raw"][order])
    for k in ("x", "y", "player", "ranks"):
        assert np.array_equal(rec[k], golden[k][order])
    # ... and as a packed corpus, read back by the native loader
    n = makedata.transcribe_to_corpus(list(texts[:3]), U.ctx(), tmp_path / "g.dgc")
    c = data.Corpus(tmp_path / "g.dgc")
    assert c.count == n == sum(len(s) for s in sels[:3])
    b = c.read(np.arange(n))
    assert np.array_equal(b["raw"], golden["raw"][np.concatenate(sels[:3])])
