"""-m gpu: board analysis on the GPU (dg_summarize_boards) -- bit-exact against the reference's own fixture corpus
(every one of the 4,398 records is an output of makedata.lua's summarize_board) and against the C oracle on random
legal positions reached by random play (captures, big groups, ladders near the edge)."""
import numpy as np
import pytest

from tests import gpu_util as U

pytestmark = pytest.mark.gpu


def test_all_reference_fixtures_bit_exact(golden):
    raw = golden["raw"]
    out = U.ctx().summarize_boards(raw[:, 0])
    ref = raw[:, [1, 2, 3, 4, 5, 7, 8]]            # liberties, liberties_after x2, kills x2, ladders x2
    bad = (out != ref).reshape(len(raw), 7, -1).any(-1)
    assert not bad.any(), (bad.sum(0), np.where(bad.any(1))[0][:10])


def _random_games(n_games, moves, seed):
    from oracle import board
    rng = np.random.default_rng(seed)
    boards = []
    for _ in range(n_games):
        s = np.zeros((19, 19), np.uint8); a = np.zeros((19, 19), np.uint8)
        player = 1
        for _ in range(moves):
            empties = np.argwhere(s == 0)
            if len(empties) == 0:
                break
            x, y = empties[rng.integers(len(empties))]
            board.update_board(s, a, int(x), int(y), player)       # captures + suicide handled as the reference does
            player = 3 - player
            if rng.random() < 0.15:
                boards.append(s.copy())
        boards.append(s.copy())
    return np.stack(boards)


def test_random_play_matches_oracle():
    from oracle import board
    boards = _random_games(12, 330, 3)               # dense boards: many captures, few liberties, ladders everywhere
    ref = board.summarize(boards)
    out = U.ctx().summarize_boards(boards)
    bad = (out != ref).reshape(len(boards), 7, -1).any(-1)
    assert not bad.any(), (bad.sum(0), np.where(bad.any(1))[0][:10])


def test_edge_cases():
    import deepgo_b200
    from oracle import board
    ctx = U.ctx()
    empty = np.zeros((1, 19, 19), np.uint8)
    out = ctx.summarize_boards(empty)
    assert np.array_equal(out, board.summarize(empty))
    assert out[0, 1, 0, 0] == 2 and out[0, 1, 9, 9] == 4 and out[0, 3:].sum() == 0     # liberties-after on the empty board
    assert ctx.summarize_boards(empty[:0]).shape == (0, 7, 19, 19)
    with pytest.raises(deepgo_b200.DeepGoError):
        ctx.summarize_boards(np.full((1, 19, 19), 3, np.uint8))


def test_update_boards_replays_every_successor(golden):
    """all 4,376 (record, next record) pairs of the corpus: stones and age after the labelled move, bit-exact"""
    raw, game, move = golden["raw"], golden["game"], golden["move"]
    idx = np.array([i for i in range(len(raw) - 1) if game[i + 1] == game[i] and move[i + 1] == move[i] + 1])
    stones = np.ascontiguousarray(raw[idx, 0]); age = np.ascontiguousarray(raw[idx, 6])
    U.ctx().update_boards(stones, age, golden["x"][idx], golden["y"][idx], golden["player"][idx])
    assert len(idx) == 4376
    assert np.array_equal(stones, raw[idx + 1, 0]) and np.array_equal(age, raw[idx + 1, 6])


def test_update_boards_suicide_and_occupied():
    import deepgo_b200
    from oracle import board
    s = np.zeros((1, 19, 19), np.uint8); a = np.zeros((1, 19, 19), np.uint8)
    s[0, 0, 1] = 1; s[0, 1, 0] = 1; a[0, 0, 1] = 254; a[0, 1, 0] = 255
    s2, a2 = s[0].copy(), a[0].copy()
    board.update_board(s2, a2, 0, 0, 2)                      # white plays the corner: suicide
    U.ctx().update_boards(s, a, [1], [1], [2])
    assert np.array_equal(s[0], s2) and np.array_equal(a[0], a2)
    assert a[0, 0, 0] == 1 and a[0, 0, 1] == 255 and a[0, 1, 0] == 255 and s[0, 0, 0] == 0
    with pytest.raises(deepgo_b200.DeepGoError, match="already been played"):
        U.ctx().update_boards(s, a, [1], [2], [2])
