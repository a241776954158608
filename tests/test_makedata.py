"""SGF front end + replay (host/makedata.py, dg_sgf_parse).  CPU: the board backend is the pinned C oracle.
The reference ships no SGF file, so the text is synthesised from the fixtures' own move lists; the records rebuilt from it must
equal the reference's records of the same games bit for bit (all 22 games, 4,398 positions)."""
import numpy as np
import pytest

import __graft_entry__ as entry

LET = "abcdefghijklmnopqrs"


class OracleBoards:
    """update_boards / summarize_boards on the CPU oracle (test infrastructure)."""

    def update_boards(self, stones, age, x, y, player):
        from oracle import board
        for i in range(len(x)):
            assert board.update_board(stones[i], age[i], int(x[i]) - 1, int(y[i]) - 1, int(player[i])) == 0

    def summarize_boards(self, stones):
        from oracle import board
        return board.summarize(stones)


def sgf_of(golden, gi, extras=True):
    sel = np.where(golden["game"] == gi)[0]
    sel = sel[np.argsort(golden["move"][sel])]
    r = golden["ranks"][sel[0]]
    s = "(;GM[1]FF[4]SZ[19]PB[a]PW[b];BR[%dd];WR[%dd];KM[5.5]" % (r[0], r[1])
    for k, i in enumerate(sel):
        s += ";%s[%s%s]" % ("BW"[golden["player"][i] - 1], LET[golden["x"][i] - 1], LET[golden["y"][i] - 1])
        if extras and k == 10:
            s += ";B[tt];W[];C[a comment]"          # passes and other properties are skipped (all_moves, makedata.lua:69-90)
    return s + ";)", sel          # (a move directly followed by ")" would be dropped by the reference's tokeniser)


@pytest.fixture(scope="module")
def md():
    entry.build()
    from deepgo_b200.host import makedata
    return makedata


def test_parser_follows_the_reference_tokenisation(md):
    p = md.parse("(;BR[3d];WR[1d];B[aa];W[sa];B[tt];W[];B[as])")
    assert p["ranks"] == (3, 1) and p["handicaps"] == []
    assert p["moves"] == [(1, 1, 1), (19, 1, 2)]                        # "B[as])" does not END with ']': the reference drops it
    assert md.parse("(;B[aa];W[sa];B[as];)")["moves"] == [(1, 1, 1), (19, 1, 2), (1, 19, 1)]
    # two properties in one piece are no token; neither is a piece followed by a bare line feed; names are not trimmed
    assert md.parse(";BR[3d]WR[1d];B[aa]")["ranks"] is None
    assert md.parse(";BR[3d];WR[1d];B[aa]\n;W[bb]; B[cc];W[dd]")["moves"] == [(2, 2, 2), (4, 4, 2)]
    # kyu ranks, missing ranks: no records for the game
    assert md.parse(";BR[3k];WR[1d];B[aa]")["ranks"] is None and md.parse(";WR[1d];B[aa]")["ranks"] is None
    assert md.parse(";BR[3d];WR[1d];WR[2d]")["ranks"] == (3, 2)           # the last one wins
    # handicap stones: only lines (CR LF!) that start with AB / AW
    p = md.parse("(;BR[3d];WR[1d]\r\nAB[pd][dp][pp]\r\nAW[dd]\r\n;W[aa];)")
    assert p["handicaps"] == [(16, 4, 1), (4, 16, 1), (16, 16, 1), (4, 4, 2)] and p["moves"] == [(1, 1, 2)]
    assert md.parse("(;BR[3d];WR[1d]\nAB[pd][dp]\n;W[aa];)")["handicaps"] == []
    import deepgo_b200
    with pytest.raises(deepgo_b200.DeepGoError):
        md.parse(";B[a!]")                                               # a column without a row: the reference crashes on it
    assert md.parse("")["moves"] == []


def test_fixture_games_rebuilt_from_sgf(md, golden):
    ngames = len(golden["games"])
    texts, sels = zip(*[sgf_of(golden, gi) for gi in range(ngames)])
    texts = list(texts) + ["(;BR[2k];WR[1d];B[aa];W[bb];)"]             # a game without dan ranks: no records
    rec = md.transcribe(texts, OracleBoards())
    order = np.concatenate(sels)
    assert len(rec["raw"]) == len(order) == 4398
    assert np.array_equal(rec["raw"], golden["raw"][order])
    for k in ("x", "y", "player", "ranks"):
        assert np.array_equal(rec[k], golden[k][order])
    assert np.array_equal(rec["game"], golden["game"][order])


def test_handicap_replay_and_corpus(md, tmp_path):
    sgf = "(;BR[5d];WR[7d]\r\nAB[dd][pp]\r\n;W[dp];B[pd];)"
    rec = md.transcribe([sgf], OracleBoards())
    assert len(rec["raw"]) == 2 and tuple(rec["ranks"][0]) == (5, 7)
    s0, a0 = rec["raw"][0, 0], rec["raw"][0, 6]
    assert s0[3, 3] == 1 and s0[15, 15] == 1 and s0.sum() == 2
    assert a0[3, 3] == 2 and a0[15, 15] == 1                           # the first handicap stone aged once under the second
    assert rec["raw"][1, 0][3, 15] == 2 and rec["raw"][1, 6][3, 15] == 1
    from deepgo_b200.host import data
    assert md.transcribe_to_corpus([sgf], OracleBoards(), tmp_path / "g.dgc") == 2
    c = data.Corpus(tmp_path / "g.dgc")
    b = c.read([0, 1])
    assert np.array_equal(b["raw"], rec["raw"]) and list(b["rank"]) == [7, 5] and list(b["player"]) == [2, 1]


def test_parser_against_the_python_restatement(md):
    """dg_sgf_parse (C++) against oracle/sgf.py on random SGF-like text built from the characters that matter."""
    import deepgo_b200
    from hypothesis import given, settings, strategies as st
    from oracle import sgf as osgf

    atoms = st.sampled_from([";", "[", "]", "\r\n", "\n", "B", "W", "A", "R", "d", "k", "a", "s", "t", "p", "3", " ", ")", "(", "BR[", "WR[", "AB[", "]["])

    @settings(max_examples=400, deadline=None)
    @given(st.lists(atoms, max_size=40).map("".join))
    def check(text):
        ranks, moves, hc = osgf.get_ranks(text), osgf.all_moves(text), osgf.handicaps(text)
        bad = any(m[1] is None for m in moves) or any(h[0] is None or h[1] is None for h in hc)
        bad = bad or (ranks is not None and any(r != int(r) or not 0 <= r <= 255 for r in ranks))
        try:
            p = md.parse(text)
        except deepgo_b200.DeepGoError:
            assert bad, text
            return
        assert not bad, text
        assert p["moves"] == moves and p["handicaps"] == hc
        assert p["ranks"] == (None if ranks is None else (int(ranks[0]), int(ranks[1])))

    check()
